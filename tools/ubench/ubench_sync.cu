// Micro-benchmark: cost of the hand-shake instructions the softmax warps execute once per key block -- mbarrier polls on a
// completed phase, arrives, TMEM load / store round trips, fences, votes -- as seen by ONE warp (clk per operation,
// back-to-back dependent repetitions).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_sync ubench_sync.cu
#include <cstdio>
#include <cstdlib>
#include "../../flash-attention-annotated_b200/csrc/fa100_ptx.cuh"
using namespace fa100;

template <int MODE>
__global__ void __launch_bounds__(128, 1) k_sync(int iters, long long* out, float* sink) {
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], (1 << 20) - 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(&tmem_base, 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tmem_base + (uint32_t(warp * 32) << 16);
  if (threadIdx.x == 0) mbar_arrive(&bar[0]);  // phase 0 of bar[0] is complete from here on
  __syncthreads();
  uint32_t acc = 0;
  uint32_t r[4][32];
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 32; ++i) r[c][i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) { acc += mbar_try_wait(&bar[0], 0); }                      // poll, phase long complete
    if (MODE == 1) { acc += mbar_test_wait(&bar[0], 0); }
    if (MODE == 2) { if (lane == 0) mbar_arrive(&bar[1]); __syncwarp(); }      // arrive (never completes a phase)
    if (MODE == 3) {                                                           // S load: 4 x 32 columns + wait
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tmem + c * 32, r[c]);
      tmem_ld_wait();
      acc += r[0][0] + r[1][1] + r[2][2] + r[3][3];
    }
    if (MODE == 4) {                                                           // P store: 4 x 16 columns + wait
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_st16(tmem + 128 + c * 16, r[c]);
      tmem_st_wait();
    }
    if (MODE == 5) { tc_fence_before(); tc_fence_after(); }
    if (MODE == 6) { acc += __any_sync(0xffffffffu, (acc + it) & 1); }
    if (MODE == 7) {                                                           // one 32-column load + wait (latency)
      tmem_ld32(tmem, r[0]); tmem_ld_wait(); acc += r[0][0];
    }
    if (MODE == 8) {                                                           // the whole per-block hand-shake, no math
      acc += mbar_try_wait(&bar[0], 0); tc_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tmem + c * 32, r[c]);
      tmem_ld_wait(); tc_fence_before(); __syncwarp();
      if (lane == 0) mbar_arrive(&bar[1]);
      acc += r[0][0] + r[1][1] + r[2][2] + r[3][3];
      acc += mbar_try_wait(&bar[0], 0); tc_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_st16(tmem + 128 + c * 16, r[c]);
      tmem_st_wait(); tc_fence_before(); __syncwarp();
      if (lane == 0) mbar_arrive(&bar[1]);
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  sink[threadIdx.x] = float(acc);
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

int main() {
  long long* d; float* sink; cudaMalloc(&d, 64); cudaMalloc(&sink, 4096);
  const int iters = 4000;
  auto report = [&](const char* name) {
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%-62s %7.1f clk per repetition  %s\n", name, double(h) / iters, cudaGetErrorString(cudaGetLastError()));
  };
  k_sync<0><<<1, 128>>>(iters, d, sink); report("mbarrier.try_wait, phase complete (4 warps polling)");
  k_sync<1><<<1, 128>>>(iters, d, sink); report("mbarrier.test_wait, phase complete");
  k_sync<2><<<1, 128>>>(iters, d, sink); report("mbarrier.arrive by lane 0 + __syncwarp");
  k_sync<3><<<1, 128>>>(iters, d, sink); report("tcgen05.ld 4 x (32x32b.x32) + wait::ld (16 KB per warp)");
  k_sync<7><<<1, 128>>>(iters, d, sink); report("tcgen05.ld 1 x (32x32b.x32) + wait::ld");
  k_sync<4><<<1, 128>>>(iters, d, sink); report("tcgen05.st 4 x (32x32b.x16) + wait::st (8 KB per warp)");
  k_sync<5><<<1, 128>>>(iters, d, sink); report("tcgen05.fence::before + ::after_thread_sync");
  k_sync<6><<<1, 128>>>(iters, d, sink); report("__any_sync (dependent)");
  k_sync<8><<<1, 128>>>(iters, d, sink); report("whole per-block hand-shake of a softmax warp, no math");
  return 0;
}
