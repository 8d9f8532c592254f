// Micro-benchmark: exp2 throughput options for the softmax warps (clk per 32-lane "warp element" per SMSP).
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cstdint>

__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t ex2_bf16x2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t ex2_f16x2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
// degree-3 polynomial 2^x for x <= 0 on the FMA/ALU pipes (Cody-Waite split with the magic-number floor)
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.0f);
  float r; asm("add.rm.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(x), "f"(12582912.0f));  // 1.5 * 2^23: floor(x) in the low mantissa bits
  const float fl = r - 12582912.0f;
  const float f = x - fl;  // [0, 1)
  float p = fmaf(f, 0.077119089663028717f, 0.227564394474029541f);
  p = fmaf(p, f, 0.695146143436431884f);
  p = fmaf(p, f, 1.0f);
  return __uint_as_float(__float_as_uint(p) + (__float_as_uint(r) << 23));
}

template <int MODE>
__global__ void __launch_bounds__(256, 1) k_exp(float* out, int iters, long long* clk) {
  float x[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) x[i] = -0.01f * (threadIdx.x + i);
  float s0 = 0.f, s1 = 0.f;
  uint32_t packed = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {  // pure MUFU
#pragma unroll
      for (int i = 0; i < 32; ++i) x[i] = ex2(x[i]) - 1.5f;
    } else if (MODE == 1) {  // softmax mix: FFMA + MUFU + FADD + pack
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float a = ex2(fmaf(x[i], 0.125f, -s0 * 1e-30f)), b = ex2(fmaf(x[i + 1], 0.125f, -s0 * 1e-30f));
        s0 += a; s1 += b;
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        packed ^= *reinterpret_cast<uint32_t*>(&v);
        x[i] = a - 1.5f; x[i + 1] = b - 1.5f;
      }
    } else if (MODE == 2) {  // bf16x2 MUFU: two elements per op
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        __nv_bfloat162 v = __floats2bfloat162_rn(x[i], x[i + 1]);
        uint32_t r = ex2_bf16x2(*reinterpret_cast<uint32_t*>(&v));
        packed ^= r;
        x[i] = __uint_as_float(r << 16) - 1.5f; x[i + 1] = __uint_as_float(r & 0xffff0000u) - 1.5f;
      }
    } else if (MODE == 3) {  // polynomial only
#pragma unroll
      for (int i = 0; i < 32; ++i) x[i] = ex2_poly(x[i]) - 1.5f;
    } else if (MODE == 4) {  // half MUFU, half polynomial, softmax mix
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float a = ex2(fmaf(x[i], 0.125f, -s0 * 1e-30f)), b = ex2_poly(fmaf(x[i + 1], 0.125f, -s0 * 1e-30f));
        s0 += a; s1 += b;
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        packed ^= *reinterpret_cast<uint32_t*>(&v);
        x[i] = a - 1.5f; x[i + 1] = b - 1.5f;
      }
    } else if (MODE == 5) {  // f16x2 MUFU
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        __half2 v = __floats2half2_rn(x[i], x[i + 1]);
        uint32_t r = ex2_f16x2(*reinterpret_cast<uint32_t*>(&v));
        packed ^= r;
        __half2 h = *reinterpret_cast<__half2*>(&r);
        x[i] = __low2float(h) - 1.5f; x[i + 1] = __high2float(h) - 1.5f;
      }
    } else if (MODE == 6) {  // 3/4 MUFU, 1/4 polynomial
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float a = ex2(fmaf(x[i], 0.125f, -s0 * 1e-30f)), b = ex2(fmaf(x[i + 1], 0.125f, -s0 * 1e-30f));
        float c = ex2(fmaf(x[i + 2], 0.125f, -s0 * 1e-30f)), d = ex2_poly(fmaf(x[i + 3], 0.125f, -s0 * 1e-30f));
        s0 += a + c; s1 += b + d;
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b), w = __floats2bfloat162_rn(c, d);
        packed ^= *reinterpret_cast<uint32_t*>(&v) ^ *reinterpret_cast<uint32_t*>(&w);
        x[i] = a - 1.5f; x[i + 1] = b - 1.5f; x[i + 2] = c - 1.5f; x[i + 3] = d - 1.5f;
      }
    }
  }
  long long t1 = clock64();
  float acc = s0 + s1 + __uint_as_float(packed & 0x3fffffffu);
#pragma unroll
  for (int i = 0; i < 32; ++i) acc += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}

template <int MODE> void run(const char* name, float* out, long long* clk) {
  for (int threads : {128, 256}) {
    const int iters = 2000;
    k_exp<MODE><<<148, threads>>>(out, iters, clk);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    const double per = double(h) / (double(iters) * 32);
    printf("%-34s warps/SMSP=%d: %.2f clk per warp-element  (%.2f clk per element per SMSP)  -> 128x128 tile on 4 SMSPs: %.0f clk\n", name, threads / 128, per,
           per / (threads / 128), per / (threads / 128) * 128 * 128 / 32 / 4);
  }
}

int main() {
  float* out; long long* clk;
  cudaMalloc(&out, 148 * 256 * 4); cudaMalloc(&clk, 8);
  run<0>("MUFU.EX2 f32 only", out, clk);
  run<1>("softmax mix (ffma+ex2+fadd+pack)", out, clk);
  run<2>("ex2.bf16x2 (+cvt)", out, clk);
  run<5>("ex2.f16x2 (+cvt)", out, clk);
  run<3>("polynomial only", out, clk);
  run<4>("mix 1/2 MUFU 1/2 poly", out, clk);
  run<6>("mix 3/4 MUFU 1/4 poly", out, clk);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
