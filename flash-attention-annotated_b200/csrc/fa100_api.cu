// fa100_api.cu -- extern "C" entry points of libfa100.so: errors, planning heuristics, tensor-map encoding.
// Role of the host half of hopper/flash_api.cpp (get_num_splits :442-470, get_pack_gqa :426-440,
// heuristics.h:9-59) -- policy re-derived for 148 SMs and 2x128-row CTA tiles, not transcribed.
#include "fa100_host.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

namespace fa100 {

static thread_local char g_err[512] = "";

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
const char* last_error() { return g_err; }

int device_sm_count(int* out) {
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  FA_CUDA(cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, dev));
  return FA100_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

int encode_tensor_map(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return set_error(FA100_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  // cuTensorMapEncodeTiled is a driver-API call: it needs the device's primary context current in THIS thread.  A thread that
  // has only inherited the device ordinal (e.g. torch's autograd engine threads, which run the backward) may not have it
  // bound yet (CUDA_ERROR_INVALID_CONTEXT); one runtime call binds it.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { FA_CUDA(cudaFree(nullptr)); ctx_bound = true; }
  CUtensorMapDataType dt = dtype == FA100_BF16   ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                           : dtype == FA100_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 1; i < rank; ++i) {
    uint64_t s = strides_bytes[i];
    // a dimension of extent 1 never moves: any legal stride will do (expanded tensors may carry 0)
    if (dims[i] == 1 && (s == 0 || (s & 15))) s = (i == 1 ? 16 : gstr[i - 2] * gdim[i - 1]);
    if (s == 0 || (s & 15))
      return set_error(FA100_ERR_INVALID, "tensor stride of dim %d (%llu bytes) must be a non-zero multiple of 16", i,
                       (unsigned long long)s);
    gstr[i - 1] = s;
  }
  if (reinterpret_cast<uintptr_t>(base) & 15) return set_error(FA100_ERR_INVALID, "tensor base must be 16-byte aligned");
  CUresult r = fn(out, dt, rank, const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(FA100_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d)", int(r), rank);
  return FA100_OK;
}

// ----------------------------------------------------------------------------- planning
static int ceil_div(int a, int b) { return (a + b - 1) / b; }

int make_plan(const Fa100Shape* s, Fa100Plan* out) {
  FA_CHECK_ARG(s && out, "null argument");
  FA_CHECK_ARG(s->h_k > 0 && s->h > 0 && s->h % s->h_k == 0, "h (%d) must be a positive multiple of h_k (%d)", s->h, s->h_k);
  std::memset(out, 0, sizeof(*out));
  int num_sm = s->num_sm;
  if (num_sm <= 0) { int rc = device_sm_count(&num_sm); if (rc) return rc; }
  const int G = s->h / s->h_k;
  const bool solo = std::max(s->d, s->dv) > 128;  // head dim > 128: one 128-row Q tile per work item (O fills TMEM)
  const int rows_cta = solo ? 128 : 256;          // else two 128-row Q tiles per CTA
  // pack GQA (rows = positions x group heads) when that fills the 256-row CTA tile >10% better, or always for
  // single-token decode where positions alone would leave 255 of 256 rows empty.
  int pack = s->pack_gqa;
  if (G == 1) pack = 0;
  else if (pack < 0) {
    auto eff = [&](int rows) { return float(rows) / float(ceil_div(rows, rows_cta) * rows_cta); };
    pack = eff(s->seqlen_q * G) > 1.1f * eff(s->seqlen_q) ? 1 : 0;
  }
  if (pack && G > 128) pack = 0;
  const int pos_per_tile = pack ? 128 / G : 128;
  const int m_blocks = std::max(1, ceil_div(s->seqlen_q, (solo ? 1 : 2) * pos_per_tile));
  const int heads_work = pack ? s->h_k : s->h;
  const int tiles = s->b * heads_work * m_blocks;
  int n_blocks = std::max(1, ceil_div(s->seqlen_k, 128));
  if ((s->is_causal || s->is_local) && s->seqlen_q > 1) n_blocks = std::max(1, std::min(n_blocks, ceil_div(s->seqlen_k, 128)));
  // Dual-stream decode: when every item's query rows fit one 128-row tile, the CTA's two tile pipelines take alternate key
  // blocks of the same rows and emit two partials per CTA-level split (fa100_fwd_roles.cuh make_geom<DUAL>), so the total
  // number of partials is even.  A caller-forced odd num_splits keeps the single-stream kernel.
  const bool dual_ok = !solo && s->seqlen_q <= pos_per_tile && n_blocks >= 2;
  int splits = s->num_splits;
  bool dual = dual_ok && (splits <= 0 || splits % 2 == 0);
  if (splits <= 0) {
    // fill the machine: smallest split count whose wave efficiency is within 85% of the best one
    splits = 1;
    const int nb_cta = dual ? ceil_div(n_blocks, 2) : n_blocks;  // key blocks one pipeline walks per unit of split
    if (tiles < 0.8f * num_sm && nb_cta > 1) {
      const int max_splits = std::min({dual ? 64 : 128, nb_cta, num_sm});
      float best = 0.f;
      auto eff = [&](int sp) {
        const float waves = float(tiles) * sp / num_sm;
        // a split only helps if it really shortens the per-CTA chain of key blocks
        const int per = ceil_div(nb_cta, sp);
        if (sp > 1 && ceil_div(nb_cta, sp - 1) == per) return 0.f;
        return waves / std::ceil(waves);
      };
      for (int sp = 1; sp <= max_splits; ++sp) best = std::max(best, eff(sp));
      for (int sp = 1; sp <= max_splits; ++sp)
        if (eff(sp) >= 0.85f * best) { splits = sp; break; }
    }
    if (dual) splits *= 2;
  }
  splits = std::max(1, std::min(splits, 128));
  const int cta_splits = dual ? splits / 2 : splits;
  out->block_m = rows_cta;
  out->block_n = 128;
  out->num_splits = splits;
  out->pack_gqa = pack;
  out->num_sm = num_sm;
  out->num_work = tiles * cta_splits;
  out->dual_stream = dual ? 1 : 0;  // dual-stream decode kernel (num_splits counts partials: two per CTA-level split)
  out->smem_bytes = 0;
  return FA100_OK;
}

}  // namespace fa100

extern "C" {

int fa100_abi_version(void) { return FA100_ABI_VERSION; }
const char* fa100_last_error(void) { return fa100::last_error(); }

int fa100_plan(const Fa100Shape* shape, Fa100Plan* out) { return fa100::make_plan(shape, out); }

static int plan_of(const Fa100FwdParams* p, Fa100Plan* plan) {
  Fa100Shape s;
  std::memset(&s, 0, sizeof(s));
  s.b = p->b; s.seqlen_q = p->seqlen_q; s.seqlen_k = p->seqlen_k; s.h = p->h; s.h_k = p->h_k; s.d = p->d; s.dv = p->dv;
  s.dtype = p->dtype; s.is_causal = p->is_causal;
  s.is_local = (p->window_left >= 0 || p->window_right >= 0 || p->attention_chunk > 0) && !p->is_causal;
  s.num_splits = p->num_splits < 1 ? 1 : p->num_splits;
  s.pack_gqa = p->pack_gqa ? 1 : 0;
  s.num_sm = p->num_sm;
  return fa100::make_plan(&s, plan);
}

int fa100_fwd(const Fa100FwdParams* p, fa100_stream_t stream) {
  if (!p) return fa100::set_error(FA100_ERR_INVALID, "null params");
  Fa100Plan plan;
  int rc = plan_of(p, &plan);
  if (rc) return rc;
  return fa100::launch_fwd(p, &plan, static_cast<cudaStream_t>(stream));
}

int64_t fa100_scheduler_metadata_words(int32_t b) { return fa100::scheduler_metadata_words(b); }

int fa100_prepare_scheduler(const Fa100FwdParams* p, int32_t* metadata, int64_t metadata_words, fa100_stream_t stream) {
  if (!p) return fa100::set_error(FA100_ERR_INVALID, "null params");
  Fa100Plan plan;
  int rc = plan_of(p, &plan);
  if (rc) return rc;
  return fa100::launch_prepare_scheduler(p, &plan, metadata, metadata_words, static_cast<cudaStream_t>(stream));
}

int fa100_combine(const Fa100CombineParams* p, fa100_stream_t stream) {
  if (!p) return fa100::set_error(FA100_ERR_INVALID, "null params");
  return fa100::launch_combine(p, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
