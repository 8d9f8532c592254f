// fa100_noalloc.cu -- split-KV forward + combine with the fp32 partials in a static device array (no allocation anywhere).
//
// No counterpart in the reference: added for callers that cannot allocate device
// memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The existing path
// (fa100_fwd with caller-supplied o_accum / lse_accum, then fa100_combine) is unchanged; this entry point runs the same two
// launches with the partials -- the only scratch memory the forward uses -- in the file-scope array below.
//
// Limits and hazards (also in include/fa100.h):
//   * the partials must fit FA100_NOALLOC_SCRATCH_BYTES (16 MiB): num_splits * b * h * seqlen_q * (dv + 1) * 4 bytes;
//     larger problems are rejected with FA100_ERR_INVALID and a message, never truncated;
//   * every GPU that loads the library holds the 16 MiB until the process exits;
//   * ALL calls share the one array and it keeps the previous call's contents: the LSE slots are reset to -inf (= "empty
//     partial") at the start of every call, and two calls that overlap on the GPU (different streams, or graphs replayed
//     concurrently) race on it -- serialise them;
//   * calls no cudaMalloc / cudaFree / allocating library routine; stream-ordered, no host synchronisation.
#include "fa100_host.h"

#include <atomic>
#include <cstring>

#include <math.h>

namespace fa100 {

constexpr size_t kNoAllocScratchBytes = size_t(FA100_NOALLOC_SCRATCH_BYTES);
static __device__ __align__(256) unsigned char g_fa100_noalloc_scratch[kNoAllocScratchBytes];

// In host code the array's name is a host-side stand-in, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, once per device, with that device current.
static float* g_scratch_ptr[128];
static std::atomic<uint64_t> g_scratch_known[2] = {{0}, {0}};

static int scratch_for_current_device(float** out) {
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  FA_CHECK_ARG(dev >= 0 && dev < 128, "device ordinal %d out of range for the no-alloc scratch table", dev);
  if (!((g_scratch_known[dev >> 6].load(std::memory_order_acquire) >> (dev & 63)) & 1)) {
    void* ptr = nullptr;
    FA_CUDA(cudaGetSymbolAddress(&ptr, g_fa100_noalloc_scratch));
    FA_CHECK_ARG(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255) == 0, "no-alloc scratch symbol has no usable device address");
    g_scratch_ptr[dev] = static_cast<float*>(ptr);  // benign race: every thread stores the same value
    g_scratch_known[dev >> 6].fetch_or(uint64_t(1) << (dev & 63), std::memory_order_release);
  }
  *out = g_scratch_ptr[dev];
  return FA100_OK;
}

// "empty partial" marker for every LSE slot: whatever the previous call left behind must not be read as a result
__global__ void fa100_noalloc_reset_lse_kernel(float* __restrict__ x, int64_t n) {
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) x[i] = -INFINITY;
}

}  // namespace fa100

extern "C" int64_t fa100_noalloc_scratch_bytes(void) { return int64_t(fa100::kNoAllocScratchBytes); }

extern "C" int fa100_fwd_splitkv_noalloc(const Fa100FwdParams* p, fa100_stream_t stream_) {
  using namespace fa100;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  FA_CHECK_ARG(p != nullptr, "null params");
  FA_CHECK_ARG(p->o != nullptr && p->lse != nullptr, "o and lse are required");
  FA_CHECK_ARG(p->o_accum == nullptr && p->lse_accum == nullptr, "o_accum / lse_accum must be null: the partials live in the library's static scratch");
  FA_CHECK_ARG(p->merge_partial == 0, "merge_partial is not available on the no-alloc path");
  FA_CHECK_ARG(p->num_splits >= 2 && p->num_splits <= 256, "num_splits must be in [2, 256] (a single pass needs no scratch: use fa100_fwd), got %d", p->num_splits);
  FA_UNSUPPORTED(p->cu_seqlens_q != nullptr, "packed (cu_seqlens_q) queries are not supported on the no-alloc path");
  FA_CHECK_ARG(p->b > 0 && p->h > 0 && p->seqlen_q > 0 && p->dv > 0 && p->dv % 4 == 0, "bad sizes b=%d h=%d seqlen_q=%d dv=%d", p->b, p->h, p->seqlen_q, p->dv);
  FA_CHECK_ARG(p->dtype == FA100_BF16 || p->dtype == FA100_FP16 || p->dtype == FA100_E4M3, "unsupported dtype %d", p->dtype);

  // the plan fa100_fwd would make (same arithmetic as fa100_api.cu plan_of)
  Fa100Shape s;
  std::memset(&s, 0, sizeof(s));
  s.b = p->b; s.seqlen_q = p->seqlen_q; s.seqlen_k = p->seqlen_k; s.h = p->h; s.h_k = p->h_k; s.d = p->d; s.dv = p->dv;
  s.dtype = p->dtype; s.is_causal = p->is_causal;
  s.is_local = (p->window_left >= 0 || p->window_right >= 0 || p->attention_chunk > 0) && !p->is_causal;
  s.num_splits = p->num_splits;
  s.pack_gqa = p->pack_gqa ? 1 : 0;
  s.num_sm = p->num_sm;
  Fa100Plan plan;
  int rc = make_plan(&s, &plan);
  if (rc) return rc;
  const int splits = plan.num_splits;

  const int64_t rows = int64_t(p->b) * p->h * p->seqlen_q;             // one partial's rows
  const int64_t o_floats = int64_t(splits) * rows * p->dv;
  const int64_t lse_floats = int64_t(splits) * rows;
  const int64_t o_floats_pad = (o_floats + 63) / 64 * 64;             // LSE block starts 256-byte aligned
  const int64_t need = (o_floats_pad + lse_floats) * 4;
  if (need > int64_t(kNoAllocScratchBytes))
    return set_error(FA100_ERR_INVALID,
                     "no-alloc scratch too small: %d splits x %d x %d x %d rows x (dv %d + 1) fp32 = %lld bytes > the static %lld bytes "
                     "(FA100_NOALLOC_SCRATCH_BYTES); use fa100_fwd with caller-supplied o_accum / lse_accum",
                     splits, p->b, p->h, p->seqlen_q, p->dv, (long long)need, (long long)kNoAllocScratchBytes);

  float* scratch = nullptr;
  rc = scratch_for_current_device(&scratch);
  if (rc) return rc;
  float* o_acc = scratch;
  float* lse_acc = scratch + o_floats_pad;

  fa100_noalloc_reset_lse_kernel<<<int((lse_floats + 255) / 256 < 592 ? (lse_floats + 255) / 256 : 592), 256, 0, stream>>>(lse_acc, lse_floats);
  FA_CUDA(cudaGetLastError());

  Fa100FwdParams fp = *p;
  fp.o_accum = o_acc;
  fp.lse_accum = lse_acc;
  fp.num_splits = splits;
  rc = launch_fwd(&fp, &plan, stream);
  if (rc) return rc;

  // partials (splits, b, h, sq, dv) / (splits, b, h, sq)  ->  o (b, sq, h, dv) in the output dtype, lse (b, h, sq)
  Fa100CombineParams c;
  std::memset(&c, 0, sizeof(c));
  c.o_partial = o_acc; c.lse_partial = lse_acc; c.o = p->o; c.lse = p->lse;
  c.op_split_stride = rows * p->dv; c.op_batch_stride = int64_t(p->h) * p->seqlen_q * p->dv; c.op_head_stride = int64_t(p->seqlen_q) * p->dv; c.op_row_stride = p->dv;
  c.lp_split_stride = rows; c.lp_batch_stride = int64_t(p->h) * p->seqlen_q; c.lp_head_stride = p->seqlen_q; c.lp_row_stride = 1;
  c.o_batch_stride = p->o_batch_stride; c.o_row_stride = p->o_row_stride; c.o_head_stride = p->o_head_stride;
  c.lse_batch_stride = int64_t(p->h) * p->seqlen_q; c.lse_head_stride = p->seqlen_q; c.lse_row_stride = 1;
  c.num_splits = splits; c.b = p->b; c.s = p->seqlen_q; c.h = p->h; c.d = p->dv;
  c.out_dtype = p->dtype == FA100_FP16 ? FA100_FP16 : FA100_BF16;
  c.seqused = p->seqused_q;  // rows past a batch element's used length stay as the caller had them
  return launch_combine(&c, stream);
}
