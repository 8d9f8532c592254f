// Micro-benchmark: one softmax tile step (TMEM load, row max, exp2, pack, TMEM store, row sum) in isolation -- no MMA,
// no barriers -- to find the instruction-level floor for 1 row/thread (128 cols) and 2 threads/row (64 cols) layouts.
#include <cstdio>
#include <cstdlib>
#include "../../flash-attention-annotated_b200/csrc/fa100_ptx.cuh"
using namespace fa100;
__device__ __forceinline__ float ex2a(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t packbf(float lo, float hi) { __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi); return *reinterpret_cast<uint32_t*>(&v); }

// exp2 of two values on the FMA/ALU pipes (Cody-Waite split by the 1.5*2^23 magic add, degree-3 polynomial, exponent
// splice), packed f32x2 arithmetic.  Inputs must be <= ~120; values below -126 are clamped.
__device__ __forceinline__ void ex2_poly2(float& a, float& b) {
  asm("{\n\t.reg .b64 x, r, f, pl;\n\t.reg .f32 a0, b0, r0, r1, p0, p1;\n\t.reg .b32 i0, i1, q0, q1;\n\t"
      "max.ftz.f32 a0, %0, 0fC2FC0000;\n\tmax.ftz.f32 b0, %1, 0fC2FC0000;\n\t"
      "mov.b64 x, {a0, b0};\n\t"
      "mov.b64 pl, {0f4B400000, 0f4B400000};\n\t"
      "add.rm.ftz.f32x2 r, x, pl;\n\t"          // floor(x) in the low mantissa bits
      "sub.rn.ftz.f32x2 f, r, pl;\n\t"           // floor(x) as a float
      "sub.rn.ftz.f32x2 f, x, f;\n\t"            // frac in [0, 1)
      "mov.b64 pl, {0f3D9DF09D, 0f3D9DF09D};\n\t" // c3 = 0.077119
      "mov.b64 x, {0f3E6906A4, 0f3E6906A4};\n\t"  // c2 = 0.227564
      "fma.rn.ftz.f32x2 pl, pl, f, x;\n\t"
      "mov.b64 x, {0f3F31F519, 0f3F31F519};\n\t"  // c1 = 0.695146
      "fma.rn.ftz.f32x2 pl, pl, f, x;\n\t"
      "mov.b64 x, {0f3F800000, 0f3F800000};\n\t"
      "fma.rn.ftz.f32x2 pl, pl, f, x;\n\t"
      "mov.b64 {p0, p1}, pl;\n\tmov.b64 {r0, r1}, r;\n\t"
      "mov.b32 i0, r0;\n\tmov.b32 i1, r1;\n\tshl.b32 i0, i0, 23;\n\tshl.b32 i1, i1, 23;\n\t"
      "mov.b32 q0, p0;\n\tmov.b32 q1, p1;\n\tadd.s32 q0, q0, i0;\n\tadd.s32 q1, q1, i1;\n\t"
      "mov.b32 %0, q0;\n\tmov.b32 %1, q1;\n\t}"
      : "+f"(a), "+f"(b));
}

// packed f32x2 helpers: (a, b) = (a, b) * s + m ;  (acc0, acc1) += (a, b)
__device__ __forceinline__ void ffma2(float& a, float& b, float sc, float m) {
  asm("{\n\t.reg .b64 x, s2, m2;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 s2, {%2, %2};\n\tmov.b64 m2, {%3, %3};\n\t"
      "fma.rn.ftz.f32x2 x, x, s2, m2;\n\tmov.b64 {%0, %1}, x;\n\t}" : "+f"(a), "+f"(b) : "f"(sc), "f"(m));
}
__device__ __forceinline__ void fadd2(float& acc0, float& acc1, float a, float b) {
  asm("{\n\t.reg .b64 x, y;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 y, {%2, %3};\n\tadd.rn.ftz.f32x2 x, x, y;\n\tmov.b64 {%0, %1}, x;\n\t}"
      : "+f"(acc0), "+f"(acc1) : "f"(a), "f"(b));
}

// VARIANT 0: 128 columns per thread, chunked (load all, max, then per chunk ffma/ex2/pack/store), sum at end
// VARIANT 1: 64 columns per thread (two chunks)
// VARIANT 2: 128 columns, per-chunk reload (two-pass)
template <int VARIANT, int NP>
__global__ void __launch_bounds__(VARIANT == 1 ? 512 : 256, 1) k_softmax(int iters, int nwarps, float scale, long long* out, float* sink) {
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x / 32;
  if (warp == 0) { tmem_alloc(&tmem_base, 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tmem_base;
  float m_run = -1e30f, l_run = 0.f;
  long long t0 = 0, t1 = 0;
  if (warp < nwarps) {
    const uint32_t tS = tmem + (uint32_t((warp & 3) * 32) << 16) + ((warp >> 2) & 1) * 128 + (VARIANT == 1 ? ((warp >> 3) & 1) * 64 : 0);
    // seed S with something finite
    { uint32_t z[32]; for (int i = 0; i < 32; ++i) z[i] = __float_as_uint(-0.01f * (i + threadIdx.x % 7)); for (int c = 0; c < 4; ++c) tmem_st32((tS & ~127u) + (tS & 128u) + c * 32, z); tmem_st_wait(); }
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      constexpr int NC = VARIANT == 1 ? 2 : 4;
      uint32_t s[NC][32];
#pragma unroll
      for (int c = 0; c < NC; ++c) tmem_ld32(tS + c * 32, s[c]);
      tmem_ld_wait();
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int c = 0; c < NC; ++c)
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(s[c][i])); mx1 = fmaxf(mx1, __uint_as_float(s[c][i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(s[c][i + 2])); mx3 = fmaxf(mx3, __uint_as_float(s[c][i + 3]));
        }
      const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      const float m_new = fmaxf(m_run, m_blk);
      const float acc = ex2a((m_run - m_new) * scale);
      m_run = m_new;
      const float m_sub = m_run * scale;
      float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
#pragma unroll
      for (int c = 0; c < NC; ++c) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {  // of every 8 elements the last NP go through the polynomial
          float x0 = __uint_as_float(s[c][i]), x1 = __uint_as_float(s[c][i + 1]);
          if (VARIANT == 3) ffma2(x0, x1, scale, -m_sub);
          else { x0 = fmaf(x0, scale, -m_sub); x1 = fmaf(x1, scale, -m_sub); }
          if ((i & 7) >= 8 - NP) ex2_poly2(x0, x1); else { x0 = ex2a(x0); x1 = ex2a(x1); }
          s[c][i] = __float_as_uint(x0); s[c][i + 1] = __float_as_uint(x1);
        }
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = packbf(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1]));
        tmem_st16(tS + c * 16, pk);
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          if (VARIANT == 3) {
            fadd2(sum0, sum1, __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
            fadd2(sum2, sum3, __uint_as_float(s[c][i + 2]), __uint_as_float(s[c][i + 3]));
          } else {
            sum0 += __uint_as_float(s[c][i]); sum1 += __uint_as_float(s[c][i + 1]);
            sum2 += __uint_as_float(s[c][i + 2]); sum3 += __uint_as_float(s[c][i + 3]);
          }
        }
      }
      tmem_st_wait();
      tc_fence_before();
      l_run = l_run * acc + ((sum0 + sum1) + (sum2 + sum3));
    }
    t1 = clock64();
  }
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = l_run + m_run;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  long long* d; float* sink; cudaMalloc(&d, 64); cudaMalloc(&sink, 148 * 512 * 4);
  const int iters = 2000;
  auto report = [&](const char* name, int nwarps, int elems_per_thread) {
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    const double per = double(h) / iters;
    printf("%-46s warps=%2d: %7.0f clk per step per warp; per SMSP per element: %.2f clk  %s\n", name, nwarps, per,
           per / (elems_per_thread * (nwarps / 4.0)), cudaGetErrorString(cudaGetLastError()));
  };
  for (int nw : {4, 8}) { k_softmax<0, 0><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread, all MUFU", nw, 128); }
  for (int nw : {4, 8}) { k_softmax<0, 2><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread, 2/8 polynomial", nw, 128); }
  for (int nw : {4, 8}) { k_softmax<3, 0><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread, all MUFU, f32x2 fma/add", nw, 128); }
  for (int nw : {4, 8}) { k_softmax<3, 2><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread, 2/8 poly, f32x2 fma/add", nw, 128); }
  for (int nw : {4, 8}) { k_softmax<0, 4><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread, 4/8 polynomial", nw, 128); }
  for (int nw : {4, 8}) { k_softmax<0, 6><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread, 6/8 polynomial", nw, 128); }
  for (int nw : {4, 8}) { k_softmax<0, 8><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread, all polynomial", nw, 128); }
  return 0;
}
