// fa100_combine.cu -- LSE-weighted merge of split-KV partials.
// Math of hopper/flash_fwd_combine_kernel.h:345-458 (LSE = max + ln sum exp(LSE_s - max),
// O = sum_s exp(LSE_s - LSE) O_s, splits with LSE = -inf skipped, summation in split order so the result is
// bitwise repeatable); layout is B200-first: HBM-bound streaming kernel, one warp per (batch, row, head)
// output row, 16-byte loads along d so every warp request is a full 512-byte line set.
#include "fa100_host.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <math.h>

namespace fa100 {

constexpr int kCombineWarps = 4;

template <int OUT>  // FA100_BF16 / FA100_FP16 / FA100_FP32
__global__ void __launch_bounds__(kCombineWarps * 32) fa100_combine_kernel(const Fa100CombineParams p) {
  asm volatile("griddepcontrol.wait;\n" ::: "memory");               // PDL: the partials come from the forward kernel in front
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t rows_total = int64_t(p.b) * p.s * p.h;
  for (int64_t r = int64_t(blockIdx.x) * kCombineWarps + warp; r < rows_total; r += int64_t(gridDim.x) * kCombineWarps) {
    const int hh = int(r % p.h);
    const int ss = int((r / p.h) % p.s);
    const int bb = int(r / (int64_t(p.h) * p.s));
    // rows past the sequence's used length are left as the caller had them (hopper/flash_fwd_combine_kernel.h, Varlen)
    if (p.seqused != nullptr && ss >= p.seqused[bb]) continue;
    const float* lp = p.lse_partial + bb * p.lp_batch_stride + ss * p.lp_row_stride + hh * p.lp_head_stride;
    // 1. max and sum over splits (lanes stride over splits)
    float mx = -INFINITY;
    for (int sp = lane; sp < p.num_splits; sp += 32) mx = fmaxf(mx, lp[sp * p.lp_split_stride]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float mx_use = mx == -INFINITY ? 0.f : mx;
    float sum = 0.f;
    for (int sp = lane; sp < p.num_splits; sp += 32) sum += expf(lp[sp * p.lp_split_stride] - mx_use);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float lse = (sum == 0.f || sum != sum) ? -INFINITY : logf(sum) + mx_use;
    const float lse_use = lse == -INFINITY ? 0.f : lse;
    // 2. weighted sum, 4 columns per lane per pass, splits in order
    const float* op = p.o_partial + bb * p.op_batch_stride + ss * p.op_row_stride + hh * p.op_head_stride;
    for (int c0 = lane * 4; c0 < p.d; c0 += 128) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int sp = 0; sp < p.num_splits; ++sp) {
        const float l = lp[sp * p.lp_split_stride];  // same address across the warp: one broadcast load
        if (l == -INFINITY) continue;
        const float wgt = expf(l - lse_use);
        const float4 v = *reinterpret_cast<const float4*>(op + sp * p.op_split_stride + c0);
        acc.x += wgt * v.x; acc.y += wgt * v.y; acc.z += wgt * v.z; acc.w += wgt * v.w;
      }
      const int64_t off = bb * p.o_batch_stride + ss * p.o_row_stride + hh * p.o_head_stride + c0;
      if (OUT == FA100_FP32) {
        *reinterpret_cast<float4*>(static_cast<float*>(p.o) + off) = acc;
      } else if (OUT == FA100_BF16) {
        __nv_bfloat162 a = __floats2bfloat162_rn(acc.x, acc.y), b2 = __floats2bfloat162_rn(acc.z, acc.w);
        uint2 u = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b2));
        *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p.o) + off) = u;
      } else {
        __half2 a = __floats2half2_rn(acc.x, acc.y), b2 = __floats2half2_rn(acc.z, acc.w);
        uint2 u = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b2));
        *reinterpret_cast<uint2*>(static_cast<__half*>(p.o) + off) = u;
      }
    }
    // the merged LSE goes out last: o / lse may alias split 0 of the partials (in-place running merge of ring attention),
    // and the weights above re-read the split LSEs
    __syncwarp();
    if (lane == 0) p.lse[bb * p.lse_batch_stride + ss * p.lse_row_stride + hh * p.lse_head_stride] = lse;
  }
}

int launch_combine(const Fa100CombineParams* p, cudaStream_t stream) {
  FA_CHECK_ARG(p->o_partial && p->lse_partial && p->o && p->lse, "null tensor");
  FA_CHECK_ARG(p->num_splits >= 1 && p->num_splits <= 256, "num_splits must be in [1, 256], got %d", p->num_splits);
  FA_CHECK_ARG(p->d % 4 == 0, "head dim must be a multiple of 4 (caller pads), got %d", p->d);
  FA_CHECK_ARG(p->op_split_stride % 4 == 0 && p->op_batch_stride % 4 == 0 && p->op_row_stride % 4 == 0 && p->op_head_stride % 4 == 0 &&
                   reinterpret_cast<uintptr_t>(p->o_partial) % 16 == 0,
               "o_partial must be 16-byte aligned with strides that are multiples of 4");
  FA_CHECK_ARG(p->o_batch_stride % 4 == 0 && p->o_row_stride % 4 == 0 && p->o_head_stride % 4 == 0, "o strides must be multiples of 4");
  const int64_t rows = int64_t(p->b) * p->s * p->h;
  if (rows == 0) return FA100_OK;
  int num_sm = 0;
  int rc = device_sm_count(&num_sm);
  if (rc) return rc;
  const int64_t blocks_needed = (rows + kCombineWarps - 1) / kCombineWarps;
  const int grid = int(blocks_needed < int64_t(num_sm) * 16 ? blocks_needed : int64_t(num_sm) * 16);
  if (p->out_dtype == FA100_FP32) FA_CUDA(launch_pdl(fa100_combine_kernel<FA100_FP32>, dim3(grid), dim3(kCombineWarps * 32), size_t(0), stream, 1, *p));
  else if (p->out_dtype == FA100_BF16) FA_CUDA(launch_pdl(fa100_combine_kernel<FA100_BF16>, dim3(grid), dim3(kCombineWarps * 32), size_t(0), stream, 1, *p));
  else if (p->out_dtype == FA100_FP16) FA_CUDA(launch_pdl(fa100_combine_kernel<FA100_FP16>, dim3(grid), dim3(kCombineWarps * 32), size_t(0), stream, 1, *p));
  else return set_error(FA100_ERR_INVALID, "bad out_dtype %d", p->out_dtype);
  FA_CUDA(cudaGetLastError());
  return FA100_OK;
}

}  // namespace fa100
