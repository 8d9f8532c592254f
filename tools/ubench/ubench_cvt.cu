// Micro-benchmark: throughput of the fp32 -> 16-bit pack (F2FP) alone and next to MUFU.EX2, and cheaper packs.
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a)); return r; }
__device__ __forceinline__ uint32_t pack_f16(float a, float b) { uint32_t r; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a)); return r; }
__device__ __forceinline__ uint32_t pack_trunc(float a, float b) { uint32_t r; asm volatile("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(__float_as_uint(a)), "r"(__float_as_uint(b))); return r; }
__device__ __forceinline__ uint32_t pack_rhu(float a, float b) {  // round half up: add 0x8000 then take the high halves
  uint32_t r; asm volatile("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(__float_as_uint(a) + 0x8000u), "r"(__float_as_uint(b) + 0x8000u)); return r; }
__device__ __forceinline__ uint16_t pack_e4m3(float a, float b) { uint16_t r; asm volatile("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(b), "f"(a)); return r; }

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(float* out, int iters, long long* clk) {
  float x[32]; uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) x[i] = -0.01f * (threadIdx.x + i);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const float m = it * 1e-7f;
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_bf16(x[i] + m, x[i + 1] + m);
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_f16(x[i] + m, x[i + 1] + m);
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_trunc(x[i] + m, x[i + 1] + m);
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_rhu(x[i] + m, x[i + 1] + m);
    } else if (MODE == 4) {  // ex2 + bf16 pack
      float e[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) e[i] = ex2(x[i] + m);
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_bf16(e[i], e[i + 1]);
    } else if (MODE == 5) {  // ex2 + rhu pack
      float e[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) e[i] = ex2(x[i] + m);
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_rhu(e[i], e[i + 1]);
    } else if (MODE == 6) {  // ex2 only
#pragma unroll
      for (int i = 0; i < 32; ++i) acc ^= __float_as_uint(ex2(x[i] + m));
    } else if (MODE == 7) {  // e4m3 pack
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_e4m3(x[i] + m, x[i + 1] + m);
    } else if (MODE == 8) {  // ex2 + e4m3 pack
      float e[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) e[i] = ex2(x[i] + m);
#pragma unroll
      for (int i = 0; i < 32; i += 2) acc ^= pack_e4m3(e[i], e[i + 1]);
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float(acc);
  if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}
template <int MODE> void run(const char* name, float* out, long long* clk) {
  for (int threads : {128, 256}) {
    const int iters = 4000;
    k<MODE><<<148, threads>>>(out, iters, clk);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    printf("%-28s warps/SMSP=%d: %.2f clk per element per SMSP\n", name, threads / 128, double(h) / (double(iters) * 32) / (threads / 128));
  }
}
int main() {
  float* out; long long* clk; cudaMalloc(&out, 148 * 256 * 4); cudaMalloc(&clk, 8);
  run<0>("fadd + cvt.bf16x2 (per elt)", out, clk);
  run<1>("fadd + cvt.f16x2", out, clk);
  run<2>("fadd + prmt trunc", out, clk);
  run<3>("fadd + iadd+prmt rhu", out, clk);
  run<7>("fadd + cvt.e4m3x2", out, clk);
  run<6>("fadd + ex2", out, clk);
  run<4>("fadd + ex2 + cvt.bf16x2", out, clk);
  run<5>("fadd + ex2 + rhu pack", out, clk);
  run<8>("fadd + ex2 + cvt.e4m3x2", out, clk);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
