// Micro-benchmark: one softmax tile step (TMEM load, row max, exp2, pack, TMEM store, row sum) in isolation -- no MMA,
// no barriers -- to find the instruction-level floor for 1 row/thread (128 cols) and 2 threads/row (64 cols) layouts.
#include <cstdio>
#include <cstdlib>
#include "../../flash-attention-annotated_b200/csrc/fa100_ptx.cuh"
using namespace fa100;
__device__ __forceinline__ float ex2a(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t packbf(float lo, float hi) { __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi); return *reinterpret_cast<uint32_t*>(&v); }

// VARIANT 0: 128 columns per thread, chunked (load all, max, then per chunk ffma/ex2/pack/store), sum at end
// VARIANT 1: 64 columns per thread (two chunks)
// VARIANT 2: 128 columns, per-chunk reload (two-pass)
template <int VARIANT>
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
        for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(ex2a(fmaf(__uint_as_float(s[c][i]), scale, -m_sub)));
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = packbf(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1]));
        tmem_st16(tS + c * 16, pk);
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          sum0 += __uint_as_float(s[c][i]); sum1 += __uint_as_float(s[c][i + 1]);
          sum2 += __uint_as_float(s[c][i + 2]); sum3 += __uint_as_float(s[c][i + 3]);
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
  for (int nw : {4, 8}) { k_softmax<0><<<148, 256>>>(iters, nw, 0.1275f, d, sink); report("128 cols/thread (1 row per thread)", nw, 128); }
  for (int nw : {8, 16}) { k_softmax<1><<<148, 512>>>(iters, nw, 0.1275f, d, sink); report("64 cols/thread (2 threads per row)", nw, 64); }
  return 0;
}
