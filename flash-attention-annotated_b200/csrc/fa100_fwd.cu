// fa100_fwd.cu -- validation, tensor maps and launch of the forward kernel.
// Role of set_params_fprop + run_mha_fwd + run_flash_fwd (hopper/flash_api.cpp:50-168,367-385,
// hopper/flash_fwd_launch_template.h:30-198) for the sm_100a kernel in fa100_fwd_softmax.cuh.
#include "fa100_host.h"
#include "fa100_fwd2.cuh"
#include "fa100_fwd_split.cuh"

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <numeric>

namespace fa100 {

#ifndef FA100_LPT_L2_MB
#define FA100_LPT_L2_MB 96  // K / V of the (batch, KV head) groups in flight under the length-ordered dynamic schedule
#endif

template <int HD, int ELT, bool FP8, bool F16, bool DUAL, bool PAIR = false>
static int launch_instance(const FwdKernelParams& kp, const CUtensorMap& tmQ, const CUtensorMap& tmK,
                           const CUtensorMap& tmV, const CUtensorMap& tmO, const CUtensorMap& tmQv, uint32_t fmt, int grid, cudaStream_t stream) {
  using L = FwdSmem<HD, ELT, PAIR>;
  auto kernel = fa100_fwd_kernel<HD, ELT, FP8, F16, DUAL, PAIR>;
  // the opt-in is per device (per context), not per process: one flag per device ordinal and kernel instance
  // (the reference sets the attribute on every launch, hopper/flash_fwd_launch_template.h:183-193)
  static std::atomic<uint64_t> attr_set[2] = {{0}, {0}};  // devices 0..127; benign race: the call is idempotent
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  const bool tracked = dev >= 0 && dev < 128;
  if (!tracked || !((attr_set[dev >> 6].load(std::memory_order_relaxed) >> (dev & 63)) & 1)) {
    FA_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes));
    if (tracked) attr_set[dev >> 6].fetch_or(uint64_t(1) << (dev & 63), std::memory_order_relaxed);
  }
  // PAIR: clusters of two CTAs (the pair shares every K / V tile, cta_group::2 MMAs; grid = 2 x pairs).  Always with PDL:
  // prologue and launch latency overlap the tail of the schedule / append-KV kernel in front (hopper/flash_fwd_launch_template.h:195-196)
  FA_CUDA(launch_pdl(kernel, dim3(grid), dim3(kThreads), size_t(L::kTotalBytes), stream, PAIR ? 2 : 1, kp, tmQ, tmK, tmV, tmO, tmQv, fmt));
  FA_CUDA(cudaGetLastError());
  return FA100_OK;
}

// The prefill mainloop of fa100_fwd2.cuh (one MMA warp, correction warpgroup; 512 threads).
template <int HD, int ELT, bool FP8, bool F16>
static int launch_instance2(const FwdKernelParams& kp, const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                            const CUtensorMap& tmO, uint32_t fmt, int grid, cudaStream_t stream) {
  using L2 = FwdSmem2<HD, ELT>;
  auto kernel = fa100_fwd2_kernel<HD, ELT, FP8, F16>;
  static std::atomic<uint64_t> attr_set[2] = {{0}, {0}};
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  const bool tracked = dev >= 0 && dev < 128;
  if (!tracked || !((attr_set[dev >> 6].load(std::memory_order_relaxed) >> (dev & 63)) & 1)) {
    FA_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L2::kTotalBytes));
    if (tracked) attr_set[dev >> 6].fetch_or(uint64_t(1) << (dev & 63), std::memory_order_relaxed);
  }
  FA_CUDA(launch_pdl(kernel, dim3(grid), dim3(kThreads2), size_t(L2::kTotalBytes), stream, 1, kp, tmQ, tmK, tmV, tmO, fmt));
  FA_CUDA(cudaGetLastError());
  return FA100_OK;
}

// The split-row prefill kernel of fa100_fwd_split.cuh (two softmax threads per query row; 640 threads).
template <int HD, int ELT, bool FP8, bool F16>
static int launch_instance_split(const FwdKernelParams& kp, const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                                 const CUtensorMap& tmO, const CUtensorMap& tmQv, uint32_t fmt, int grid, cudaStream_t stream) {
  using LS = FwdSmemSplit<HD, ELT>;
  auto kernel = fa100_fwd_split_kernel<HD, ELT, FP8, F16>;
  static std::atomic<uint64_t> attr_set[2] = {{0}, {0}};
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  const bool tracked = dev >= 0 && dev < 128;
  if (!tracked || !((attr_set[dev >> 6].load(std::memory_order_relaxed) >> (dev & 63)) & 1)) {
    FA_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LS::kTotalBytes));
    if (tracked) attr_set[dev >> 6].fetch_or(uint64_t(1) << (dev & 63), std::memory_order_relaxed);
  }
  FA_CUDA(launch_pdl(kernel, dim3(grid), dim3(HD > 128 ? kThreads : kSplitThreads), size_t(LS::kTotalBytes), stream, 1, kp, tmQ, tmK, tmV, tmO, tmQv, fmt));
  FA_CUDA(cudaGetLastError());
  return FA100_OK;
}

// The 2-CTA variant (cta_group::2; the second MMA form SURVEY 8(a) a2 names): bf16 / fp16, head dims of exactly 128,
// dense or packed K/V (not paged), not the dual-stream decode kernel.  Parity-green on the whole GPU suite when forced on
// and bitwise equal to the 1-CTA kernel.  Interleaved bursts in one process (tools/pair_ab.py, profiles/r01c_pair_ab.log):
// within +-1 % of the 1-CTA kernel without a mask (1269 vs 1274 and 1290 vs 1283 TFLOP/s), 1 to 4 % behind it under a causal
// mask (1208 vs 1220 MHA s32768, 1138 vs 1191 on C2: a causal pair walks the union of its two tiles' key ranges, one fully
// masked block per tile pair, and its 512-row work items balance worse).  The tensor
// chain is shorter with pairs, but the softmax chain is co-critical, so nothing is gained yet and the
// default stays one CTA per work item.  Fa100FwdParams.variant: 0 / 2 = one CTA per work item, 1 = pairs where eligible.
static bool use_pair(const Fa100FwdParams* p, const Fa100Plan* plan) {
  return (p->variant & 3) == 1 && p->dtype != FA100_E4M3 && p->d == 128 && p->dv == 128 && p->page_table == nullptr && !plan->dual_stream &&
         plan->num_sm >= 2;
}

// Everything the kernels need that is not a tensor map; shared by the forward launch and the schedule kernel.
static void fill_kernel_params(const Fa100FwdParams* p, const Fa100Plan* plan, FwdKernelParams& kp) {
  const bool fp8 = p->dtype == FA100_E4M3;
  const int elt = fp8 ? 1 : 2;
  const bool varlen_q = p->cu_seqlens_q != nullptr;
  const bool varlen_k = p->cu_seqlens_k != nullptr;
  const bool paged = p->page_table != nullptr;
  const int splits = plan->num_splits;
  std::memset(&kp, 0, sizeof(kp));
  kp.o = p->o; kp.lse = p->lse; kp.o_accum = p->o_accum; kp.lse_accum = p->lse_accum;
  kp.o_batch_stride = p->o_batch_stride; kp.o_row_stride = p->o_row_stride; kp.o_head_stride = p->o_head_stride;
  const int64_t rows_q = varlen_q ? p->total_q : p->seqlen_q;
  kp.lse_head_stride = rows_q;
  kp.lse_batch_stride = varlen_q ? 0 : int64_t(p->h) * rows_q;
  kp.lseacc_split_stride = (varlen_q ? 1 : int64_t(p->b)) * p->h * rows_q;
  kp.oacc_head_stride = rows_q * p->dv;
  kp.oacc_batch_stride = varlen_q ? 0 : int64_t(p->h) * rows_q * p->dv;
  kp.oacc_split_stride = kp.lseacc_split_stride * p->dv;
  kp.cu_seqlens_q = p->cu_seqlens_q; kp.cu_seqlens_k = p->cu_seqlens_k;
  kp.seqused_q = p->seqused_q; kp.seqused_k = p->seqused_k;
  kp.leftpad_k = p->leftpad_k; kp.kv_batch_idx = p->kv_batch_idx;
  kp.page_table = p->page_table; kp.page_table_batch_stride = p->page_table_batch_stride;
  kp.q_descale = p->q_descale; kp.k_descale = p->k_descale; kp.v_descale = p->v_descale;
  kp.q_descale_batch_stride = p->q_descale_batch_stride; kp.q_descale_head_stride = p->q_descale_head_stride;
  kp.k_descale_batch_stride = p->k_descale_batch_stride; kp.k_descale_head_stride = p->k_descale_head_stride;
  kp.v_descale_batch_stride = p->v_descale_batch_stride; kp.v_descale_head_stride = p->v_descale_head_stride;
  kp.b = p->b; kp.h = p->h; kp.h_k = p->h_k; kp.seqlen_q = p->seqlen_q; kp.seqlen_k = p->seqlen_k; kp.d = p->d; kp.dv = p->dv;
  kp.page_size = paged ? p->page_size : 0;
  // paged K/V: one TMA box never crosses a page.  gcd(page_size, 128) rows per box serves every page size (the reference
  // falls back to cp.async gathers when the page does not hold whole key blocks, hopper/paged_kv.h:233-283); pages of
  // 128 rows or a multiple are one box per key block, a page of 1 row is 128 boxes.
  kp.kv_box_rows = paged ? std::gcd(128, p->page_size) : 128;
  kp.merge = p->merge_partial;
  const int G = p->h / p->h_k;
  kp.qheads_per_kv = G;
  kp.pack = plan->pack_gqa ? G : 1;
  kp.pos_per_tile = 128 / kp.pack;
  kp.qk_boxes = (p->d * elt + 127) / 128;
  kp.v_boxes = (p->dv * elt + 127) / 128;
  kp.qk_slices = (p->d * elt + 31) / 32;
  kp.pv_n = (p->dv + 15) / 16 * 16;
  kp.qv_slices = p->qv ? (p->dv + 15) / 16 : 0;
  const bool dual = plan->dual_stream != 0;  // two partials per CTA-level split (make_geom<DUAL>)
  const int cta_splits = dual ? splits / 2 : splits;
  const bool solo = std::max(p->d, p->dv) > 128;  // one Q tile per work item (make_geom<.., SOLO>)
  const bool pair = use_pair(p, plan);            // four Q tiles per work item, two per CTA of a pair (make_geom<.., PAIR>)
  const int tiles_item = (dual || solo) ? 1 : (pair ? 4 : 2);
  kp.num_m_blocks = dual ? 1 : std::max(1, (p->seqlen_q + tiles_item * kp.pos_per_tile - 1) / (tiles_item * kp.pos_per_tile));
  kp.heads_work = plan->pack_gqa ? p->h_k : p->h;
  kp.num_splits = cta_splits;
  kp.num_work = kp.num_m_blocks * p->b * kp.heads_work * cta_splits;
  kp.tiles_per_item = tiles_item;
  kp.num_sm = plan->num_sm;
  kp.sched = p->scheduler_metadata ? p->scheduler_metadata + kSchedHeaderWords : nullptr;
  // causal / local normalisation (hopper/flash_api.cpp:148-160, 796-805)
  int wl = p->window_left, wr = p->window_right;
  bool causal = p->is_causal != 0;
  if (causal) wr = 0;
  const bool fixed_len = !varlen_q && !varlen_k && !p->seqused_q && !p->seqused_k && !p->leftpad_k;
  if (fixed_len && wl >= p->seqlen_k - 1) wl = -1;
  if (fixed_len && wr >= p->seqlen_q - 1) wr = -1;
  kp.is_causal = (wl < 0 && wr == 0 && p->attention_chunk == 0) ? 1 : 0;
  kp.is_local = ((wl >= 0 || wr >= 0 || p->attention_chunk > 0) && !kp.is_causal) ? 1 : 0;
  kp.window_left = wl; kp.window_right = wr; kp.attention_chunk = p->attention_chunk;
  kp.out_fp16 = p->dtype == FA100_FP16;
  // Dynamic scheduling over a length-ordered list (decode_work, geom_role) under causal / local masks, when the caller supplies a
  // zeroed counter: not for CTA pairs (both CTAs must agree on the item).  The
  // super-groups of the list are sized so that the K / V of the groups in flight stay in L2 (FA100_LPT_L2_MB).
  kp.tile_counter = nullptr;
  kp.lpt_groups = 0;
  if (p->tile_counter && !pair && kp.sched != nullptr) {
    // ragged batch: the device-built list is already ordered (batch elements with the most key blocks first); the CTAs take its
    // items as they finish instead of in a fixed deal
    kp.tile_counter = p->tile_counter;
  } else if (p->tile_counter && !pair && (kp.is_causal || kp.is_local) && kp.num_m_blocks > 1) {
    const int gsz = kp.heads_work / p->h_k;
    const int64_t kv_bytes = int64_t(p->seqlen_k) * (p->d + p->dv) * elt;
    // MHA (one query head per K / V head and row-block level) needs more groups in flight for an even deal and gets the whole budget;
    // with several query heads per K / V head half of it balances as well and keeps C2's DRAM traffic at the algorithmic 1.4 GB
    // (2.7 GB with the whole budget, same speed: profiles/r02p_c2_fwd_ncu_full_lpt96.md)
    const int64_t l2_bytes = (int64_t(FA100_LPT_L2_MB) << 20) / (gsz == 1 ? 1 : 2);
    const int64_t by_l2 = std::max<int64_t>(1, l2_bytes / std::max<int64_t>(1, kv_bytes));
    const int by_grid = std::max(1, plan->num_sm / std::max(1, gsz * cta_splits));
    kp.lpt_groups = int(std::min<int64_t>(std::min<int64_t>(by_l2, by_grid), int64_t(p->b) * p->h_k));
    kp.tile_counter = p->tile_counter;
  }
  kp.softmax_scale = p->softmax_scale;
  if (p->softcap > 0.f) { kp.softcap_pre = p->softmax_scale / p->softcap; kp.scale_log2 = p->softcap * kLog2e; }
  else { kp.softcap_pre = 0.f; kp.scale_log2 = p->softmax_scale * kLog2e; }

  (void)varlen_k;
}

// ----------------------------------------------------------------------------- schedule for ragged batches
// Role of prepare_varlen_num_blocks_kernel (hopper/flash_prepare_scheduler.cu:43-215): one CTA looks at every batch
// element's real lengths, counts its row blocks and key blocks with the same arithmetic the forward kernel uses
// (load_seq, tile_range), orders the batch elements by key-block count (longest first, the reference's LPT sort
// `:150-190`) and writes the prefix of work items over that order.  The forward kernel then walks a list without
// entries for rows past a sequence's end, with the expensive sequences at the front of every CTA's walk.
__global__ void __launch_bounds__(1024, 1)
fa100_schedule_kernel(const FwdKernelParams p, int32_t* __restrict__ meta, int32_t* __restrict__ scratch, float* __restrict__ fill, int64_t fill_n) {
  pdl_wait();               // the lengths may come from the kernel in front (append-KV, the caller's producer)
  pdl_launch_dependents();  // the forward's prologue may start; it waits for this grid before reading the schedule
  if (blockIdx.x > 0) {
    // CTAs 1.. : with per-batch split counts some partials are never written -- make every LSE slot read as an empty
    // partial (-inf) for the combine pass (was a separate launch)
    for (int64_t i = int64_t(blockIdx.x - 1) * blockDim.x + threadIdx.x; i < fill_n; i += int64_t(gridDim.x - 1) * blockDim.x) fill[i] = -INFINITY;
    return;
  }
  int32_t* order = meta + 2;
  int32_t* prefix = order + p.b;
  int32_t* nsplit = prefix + p.b + 1;  // per ordered batch element
  int32_t* cost = scratch;             // key blocks of the batch element (sort key)
  int32_t* mblk = scratch + p.b;       // row blocks of the batch element
  __shared__ int32_t run_total[1024];
  __shared__ unsigned long long total_blocks;
  if (threadIdx.x == 0) total_blocks = 0ull;
  __syncthreads();
  const int rows_item = p.tiles_per_item * p.pos_per_tile;
  for (int bi = threadIdx.x; bi < p.b; bi += blockDim.x) {
    const SeqInfo sq = load_seq(p, bi);
    const int m_blocks = sq.sq > 0 ? (sq.sq + rows_item - 1) / rows_item : 0;
    int c = 0;
    for (int t = 0; t < m_blocks * p.tiles_per_item; ++t) {
      const TileRange r = tile_range(p, sq, t * p.pos_per_tile, 0, 1);  // the whole range, before it is cut into splits
      c += r.n_max - r.n_min;
    }
    cost[bi] = c;
    mblk[bi] = m_blocks;
    atomicAdd(&total_blocks, (unsigned long long)c * p.heads_work);
  }
  __syncthreads();
  // Split counts per batch element (hopper/flash_prepare_scheduler.cu:96-120): every CTA-level split should hold about
  // total key blocks * 1.1 / SMs key blocks, never more splits than the caller's buffers hold (num_splits).
  const long long per_cta = max(1ll, (long long)((double)total_blocks * 1.1 / max(p.num_sm, 1)) + 1);
  // rank sort (stable): position = batches with a larger cost, or the same cost and a smaller index
  for (int bi = threadIdx.x; bi < p.b; bi += blockDim.x) {
    const int c = cost[bi];
    int rank = 0;
    for (int j = 0; j < p.b; ++j) { const int cj = cost[j]; rank += (cj > c) || (cj == c && j < bi); }
    order[rank] = bi;
    int ns = 1;
    if (p.num_splits > 1 && mblk[bi] > 0) {
      const long long blocks_item = c / (mblk[bi] * p.tiles_per_item);  // key blocks one row block walks (average)
      ns = int(min((long long)p.num_splits, max(1ll, (blocks_item + per_cta - 1) / per_cta)));
    }
    nsplit[rank] = ns;
  }
  __syncthreads();
  // exclusive prefix of work items over the order: per-thread runs, then a scan of the 1024 run totals
  const int per = (p.b + blockDim.x - 1) / blockDim.x;
  const int lo = min(int(threadIdx.x) * per, p.b), hi = min(lo + per, p.b);
  int sum = 0;
  for (int r = lo; r < hi; ++r) sum += mblk[order[r]] * p.heads_work * nsplit[r];
  run_total[threadIdx.x] = sum;
  __syncthreads();
  for (int off = 1; off < int(blockDim.x); off <<= 1) {
    const int v = threadIdx.x >= off ? run_total[threadIdx.x - off] : 0;
    __syncthreads();
    run_total[threadIdx.x] += v;
    __syncthreads();
  }
  int acc = run_total[threadIdx.x] - sum;
  for (int r = lo; r < hi; ++r) { prefix[r] = acc; acc += mblk[order[r]] * p.heads_work * nsplit[r]; }
  if (threadIdx.x == blockDim.x - 1) {
    prefix[p.b] = run_total[threadIdx.x]; meta[0] = run_total[threadIdx.x]; meta[1] = p.b;
    // the plan this schedule was built for; the forward kernel checks it (sched_matches)
    int32_t* header = meta - kSchedHeaderWords;
    header[0] = kSchedMagic; header[3] = p.num_sm; header[5] = p.tiles_per_item; header[6] = p.num_splits; header[7] = p.pack;
  }
}

// splits a batch element does not use must read as empty in the combine pass
__global__ void fa100_fill_neg_inf_kernel(float* __restrict__ x, int64_t n) {
  pdl_wait();
  pdl_launch_dependents();
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) x[i] = -INFINITY;
}

int64_t scheduler_metadata_words(int b) { return kSchedHeaderWords + 2 + int64_t(b) + (b + 1) + int64_t(b) + 2 * int64_t(b); }

int launch_prepare_scheduler(const Fa100FwdParams* p, const Fa100Plan* plan, int32_t* metadata, int64_t words, cudaStream_t stream) {
  FA_CHECK_ARG(metadata != nullptr, "metadata must be non-null");
  FA_CHECK_ARG(p->b > 0 && p->h > 0 && p->h_k > 0 && p->h % p->h_k == 0, "bad batch/head counts b=%d h=%d h_k=%d", p->b, p->h, p->h_k);
  FA_CHECK_ARG(words >= scheduler_metadata_words(p->b), "metadata holds %lld words, %lld needed for batch %d", (long long)words,
               (long long)scheduler_metadata_words(p->b), p->b);
  FwdKernelParams kp;
  fill_kernel_params(p, plan, kp);
  int32_t* body = metadata + kSchedHeaderWords;
  // the LSE partials of a split call are pre-filled here (Fa100FwdParams.lse_accum given): the forward then skips its own fill
  float* fill = nullptr;
  int64_t fill_n = 0;
  if (p->lse_accum != nullptr && plan->num_splits > 1 && p->merge_partial == 0) { fill = p->lse_accum; fill_n = kp.lseacc_split_stride * plan->num_splits; }
  const int fill_ctas = fill_n > 0 ? int(std::min<int64_t>((fill_n + 4095) / 4096, 147)) : 0;
  FA_CUDA(launch_pdl(fa100_schedule_kernel, dim3(1 + fill_ctas), dim3(1024), size_t(0), stream, 1, kp, body, body + 2 + p->b + (p->b + 1) + p->b, fill, fill_n));
  FA_CUDA(cudaGetLastError());
  return FA100_OK;
}

int launch_fwd(const Fa100FwdParams* p, const Fa100Plan* plan, cudaStream_t stream) {
  // ---- validation (hopper/flash_api.cpp:710-856)
  FA_CHECK_ARG(p->q && p->k && p->v, "q, k and v must be non-null");
  FA_CHECK_ARG(p->dtype == FA100_BF16 || p->dtype == FA100_FP16 || p->dtype == FA100_E4M3, "unsupported dtype %d", p->dtype);
  FA_CHECK_ARG(p->b > 0 && p->h > 0 && p->h_k > 0 && p->h % p->h_k == 0, "bad batch/head counts b=%d h=%d h_k=%d", p->b, p->h, p->h_k);
  FA_CHECK_ARG(p->d > 0 && p->dv > 0 && p->d % 8 == 0 && p->dv % 8 == 0, "head dims must be multiples of 8 (d=%d dv=%d)", p->d, p->dv);
  const bool fp8 = p->dtype == FA100_E4M3;
  const int elt = fp8 ? 1 : 2;
  if (fp8) FA_CHECK_ARG(p->d % 16 == 0 && p->dv % 16 == 0, "fp8 head dims must be multiples of 16");
  FA_UNSUPPORTED(p->d > 256 || p->dv > 256, "head dim > 256 is not supported (d=%d dv=%d)", p->d, p->dv);
  FA_UNSUPPORTED(p->k_new || p->v_new || p->rotary_cos, "append-KV / rotary handled by the caller-side kernel, not fa100_fwd");
  if (p->qv) {  // hopper/flash_api.cpp:1028-1050
    FA_CHECK_ARG(p->d <= 64, "q_v is only supported for head_size <= 64");
    FA_CHECK_ARG(p->dv >= 256, "q_v is only supported for hdim_v >= 256.");
    FA_CHECK_ARG(!fp8, "q_v is only supported for fp16 and bf16 data type");
    FA_UNSUPPORTED(p->dv > 256, "q_v with hdim_v > 256 is not supported (hdim_v=%d)", p->dv);
  }
  const bool varlen_q = p->cu_seqlens_q != nullptr;
  const bool varlen_k = p->cu_seqlens_k != nullptr;
  const bool paged = p->page_table != nullptr;
  FA_CHECK_ARG(!(paged && varlen_k), "page_table and cu_seqlens_k are exclusive");
  if (paged) {
    FA_CHECK_ARG(p->page_size > 0 && p->num_pages > 0, "paged KV needs page_size and num_pages");
    FA_UNSUPPORTED(p->leftpad_k != nullptr, "leftpad_k with paged KV not supported");
  }
  const int splits = plan->num_splits;
  if (splits > 1) FA_CHECK_ARG(p->o_accum && p->lse_accum, "num_splits > 1 needs o_accum and lse_accum");
  FA_CHECK_ARG(p->merge_partial >= 0 && p->merge_partial <= 2, "merge_partial must be 0, 1 or 2");
  if (p->merge_partial) FA_CHECK_ARG(p->o_accum && p->lse_accum && splits == 1 && !plan->dual_stream, "merge_partial needs o_accum / lse_accum and a single partial");
  if (p->merge_partial == 2) FA_CHECK_ARG(p->o && p->lse, "merge_partial = 2 writes o and lse");
  FA_CHECK_ARG((p->o_accum == nullptr) == (p->lse_accum == nullptr), "o_accum and lse_accum must be given together");
  FA_CHECK_ARG(p->o_accum != nullptr || (p->o != nullptr && p->lse != nullptr), "o and lse are required without o_accum");

  // ---- kernel parameters
  FwdKernelParams kp;
  fill_kernel_params(p, plan, kp);
  const bool dual = plan->dual_stream != 0;
  const bool solo = std::max(p->d, p->dv) > 128;
  const bool pair = use_pair(p, plan);
  const int64_t rows_q = varlen_q ? p->total_q : p->seqlen_q;

  // ---- tensor maps
  const int cols_per_box = 128 / elt;
  CUtensorMap tmQ, tmK, tmV, tmQv;
  {
    const uint64_t heads3 = plan->pack_gqa ? p->h_k : p->h;
    uint64_t dims[5] = {uint64_t(p->d), uint64_t(kp.pack), uint64_t(rows_q), heads3, uint64_t(varlen_q ? 1 : p->b)};
    uint64_t str[5] = {0, uint64_t(p->q_head_stride) * elt, uint64_t(p->q_row_stride) * elt,
                       uint64_t(p->q_head_stride) * kp.pack * elt, uint64_t(p->q_batch_stride) * elt};
    uint32_t box[5] = {uint32_t(cols_per_box), uint32_t(kp.pack), uint32_t(kp.pos_per_tile), 1, 1};
    int rc = encode_tensor_map(&tmQ, p->dtype, 5, p->q, dims, str, box);
    if (rc) return rc;
    tmQv = tmQ;
    if (p->qv) {  // same rows / heads / packing as q, dv columns
      uint64_t dimsv[5] = {uint64_t(p->dv), uint64_t(kp.pack), uint64_t(rows_q), heads3, uint64_t(varlen_q ? 1 : p->b)};
      uint64_t strv[5] = {0, uint64_t(p->qv_head_stride) * elt, uint64_t(p->qv_row_stride) * elt,
                          uint64_t(p->qv_head_stride) * kp.pack * elt, uint64_t(p->qv_batch_stride) * elt};
      rc = encode_tensor_map(&tmQv, p->dtype, 5, p->qv, dimsv, strv, box);
      if (rc) return rc;
    }
  }
  {
    const uint64_t rows_k = paged ? uint64_t(p->page_size) : uint64_t(varlen_k ? p->total_k : p->seqlen_k);
    const uint64_t nb = paged ? uint64_t(p->num_pages) : uint64_t(varlen_k ? 1 : p->b_k);
    uint32_t box[4] = {uint32_t(cols_per_box), uint32_t(kp.kv_box_rows), 1, 1};
    uint32_t boxK[4] = {uint32_t(cols_per_box), uint32_t(pair ? 64 : kp.kv_box_rows), 1, 1};  // a CTA of a pair holds 64 of the 128 keys
    uint64_t dimsK[4] = {uint64_t(p->d), rows_k, uint64_t(p->h_k), nb};
    uint64_t strK[4] = {0, uint64_t(p->k_row_stride) * elt, uint64_t(p->k_head_stride) * elt, uint64_t(p->k_batch_stride) * elt};
    int rc = encode_tensor_map(&tmK, p->dtype, 4, p->k, dimsK, strK, boxK);
    if (rc) return rc;
    uint64_t dimsV[4] = {uint64_t(p->dv), rows_k, uint64_t(p->h_k), nb};
    uint64_t strV[4] = {0, uint64_t(p->v_row_stride) * elt, uint64_t(p->v_head_stride) * elt, uint64_t(p->v_batch_stride) * elt};
    rc = encode_tensor_map(&tmV, p->dtype, 4, p->v, dimsV, strV, box);
    if (rc) return rc;
  }
  // O through TMA stores when a tile's rows are one dense run of one (batch, head): no packed-varlen q (a box would run
  // into the next sequence), no seqused_q, no pack-GQA (rows interleave heads), no fp32 partials; hdim <= 128.
  CUtensorMap tmO = tmQ;
  {
    // packed / seqused_q batches take the TMA path per tile (o_tma_check); the opt-in variant kernels keep it for dense batches only
    const bool partial_rows = varlen_q || p->seqused_q != nullptr;
    const bool variant_kernel = (p->variant & (FA100_VARIANT_FOUR_ROLE | FA100_VARIANT_SPLIT_ROWS)) != 0;
    const bool o_tma = !(p->variant & FA100_VARIANT_NO_TMA_STORE) && p->o_accum == nullptr && !(partial_rows && variant_kernel) && kp.pack == 1 &&
                       std::max(p->d, p->dv) <= 128 && (p->o_row_stride * 2) % 16 == 0 && (p->o_head_stride * 2) % 16 == 0 &&
                       (varlen_q || (p->o_batch_stride * 2) % 16 == 0) && reinterpret_cast<uintptr_t>(p->o) % 16 == 0;
    if (o_tma) {
      // packed: one "batch" of total_q rows (the kernel adds the sequence's row offset)
      uint64_t dims[4] = {uint64_t(p->dv), uint64_t(varlen_q ? p->total_q : p->seqlen_q), uint64_t(p->h), uint64_t(varlen_q ? 1 : p->b)};
      uint64_t str[4] = {0, uint64_t(p->o_row_stride) * 2, uint64_t(p->o_head_stride) * 2,
                         varlen_q ? uint64_t(p->o_row_stride) * 2 * uint64_t(p->total_q) : uint64_t(p->o_batch_stride) * 2};
      uint32_t box[4] = {64, 128, 1, 1};
      int rc = encode_tensor_map(&tmO, p->dtype == FA100_FP16 ? FA100_FP16 : FA100_BF16, 4, p->o, dims, str, box);
      if (rc) return rc;
      kp.o_tma = 1;
      kp.o_tma_check = partial_rows ? 1 : 0;
    }
  }
  if (p->o_accum == nullptr) {
    FA_CHECK_ARG((p->o_row_stride % 8 == 0) && (p->o_head_stride % 8 == 0) && (p->o_batch_stride % 8 == 0) &&
                     (reinterpret_cast<uintptr_t>(p->o) % 16 == 0),
                 "o must be 16-byte aligned with strides that are multiples of 8 elements");
  }

  if (kp.sched && splits > 1 && !p->lse_accum_prefilled) {
    // with per-batch split counts some partials are never written: make them read as empty (LSE = -inf) for fa100_combine.
    // fa100_prepare_scheduler does this itself when it is given lse_accum (the caller then sets lse_accum_prefilled); this
    // launch remains for a schedule built ahead of time (get_scheduler_metadata)
    const int64_t n = kp.lseacc_split_stride * splits;
    FA_CUDA(launch_pdl(fa100_fill_neg_inf_kernel, dim3(int(std::min<int64_t>((n + 255) / 256, 1184))), dim3(256), size_t(0), stream, 1, p->lse_accum, n));
    FA_CUDA(cudaGetLastError());
  }
  int num_sm = plan->num_sm;
  const int grid = pair ? 2 * std::max(1, std::min(kp.num_work, num_sm / 2)) : std::max(1, std::min(kp.num_work, num_sm));
  const uint32_t fmt = p->dtype == FA100_BF16 ? kFmtBF16 : (p->dtype == FA100_FP16 ? kFmtF16 : kFmtE4M3);
  const int hd = solo ? 256 : ((p->d <= 64 && p->dv <= 64 && !fp8) ? 64 : 128);
  const bool f16 = p->dtype == FA100_FP16;
#define FA_LAUNCH(HD_, ELT_, FP8_, F16_)                                                                      \
  (dual ? launch_instance<HD_, ELT_, FP8_, F16_, true>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream)                 \
        : launch_instance<HD_, ELT_, FP8_, F16_, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream))
  if (fp8 && hd == 256) return launch_instance<256, 1, true, false, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
  // opt-in (Fa100FwdParams.variant & FA100_VARIANT_FOUR_ROLE): the four-role prefill mainloop of fa100_fwd2.cuh for two Q tiles
  // per work item, head dims <= 128, one CTA per item
  if (!dual && !solo && !pair && (p->variant & FA100_VARIANT_FOUR_ROLE)) {
    if (fp8) return launch_instance2<128, 1, true, false>(kp, tmQ, tmK, tmV, tmO, fmt, grid, stream);
    if (hd == 64) return f16 ? launch_instance2<64, 2, false, true>(kp, tmQ, tmK, tmV, tmO, fmt, grid, stream)
                             : launch_instance2<64, 2, false, false>(kp, tmQ, tmK, tmV, tmO, fmt, grid, stream);
    return f16 ? launch_instance2<128, 2, false, true>(kp, tmQ, tmK, tmV, tmO, fmt, grid, stream)
               : launch_instance2<128, 2, false, false>(kp, tmQ, tmK, tmV, tmO, fmt, grid, stream);
  }
  if (!dual && solo && (p->variant & FA100_VARIANT_SPLIT_ROWS)) {
    if (fp8) return launch_instance_split<256, 1, true, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
    return f16 ? launch_instance_split<256, 2, false, true>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream)
               : launch_instance_split<256, 2, false, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
  }
  if (!dual && !solo && !pair && (p->variant & FA100_VARIANT_SPLIT_ROWS)) {
    if (fp8) return launch_instance_split<128, 1, true, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
    if (hd == 64) return f16 ? launch_instance_split<64, 2, false, true>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream)
                             : launch_instance_split<64, 2, false, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
    return f16 ? launch_instance_split<128, 2, false, true>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream)
               : launch_instance_split<128, 2, false, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
  }
  if (fp8) return FA_LAUNCH(128, 1, true, false);
  if (hd == 256)
    return f16 ? launch_instance<256, 2, false, true, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream)
               : launch_instance<256, 2, false, false, false>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
  if (hd == 64) return f16 ? FA_LAUNCH(64, 2, false, true) : FA_LAUNCH(64, 2, false, false);
  if (pair)
    return f16 ? launch_instance<128, 2, false, true, false, true>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream)
               : launch_instance<128, 2, false, false, false, true>(kp, tmQ, tmK, tmV, tmO, tmQv, fmt, grid, stream);
  return f16 ? FA_LAUNCH(128, 2, false, true) : FA_LAUNCH(128, 2, false, false);
#undef FA_LAUNCH
}

}  // namespace fa100

extern "C" int fa100_debug_hang_info(uint32_t out[4]) {
  // reads and clears the watchdog record of this library's forward kernels
  unsigned int h[4] = {0, 0, 0, 0}, zero[4] = {0, 0, 0, 0}, z = 0;
  cudaError_t e = cudaMemcpyFromSymbol(h, fa100::g_fa100_hang_info, sizeof(h));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(fa100::g_fa100_hang_info, zero, sizeof(zero));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(fa100::g_fa100_abort, &z, sizeof(z));
  for (int i = 0; i < 4; ++i) out[i] = h[i];
  return e == cudaSuccess ? 0 : fa100::set_error(FA100_ERR_CUDA, "hang info: %s", cudaGetErrorString(e));
}

extern "C" int fa100_debug_trace(uint32_t* out, int max_events) {
  // copies up to max_events {tag|warp<<16, begin, end} triples of the opt-in timeline and clears it;
  // returns the number of events (0 in the product build, which carries no trace code)
#ifdef FA100_TRACE
  static unsigned int host_buf[fa100::kTraceWarps * fa100::kTracePerWarp * 3];
  unsigned int cnt[fa100::kTraceWarps], zero[fa100::kTraceWarps] = {};
  if (cudaMemcpyFromSymbol(cnt, fa100::g_fa100_trace_cnt, sizeof(cnt)) != cudaSuccess) return -3;
  if (cudaMemcpyFromSymbol(host_buf, fa100::g_fa100_trace, sizeof(host_buf)) != cudaSuccess) return -3;
  cudaMemcpyToSymbol(fa100::g_fa100_trace_cnt, zero, sizeof(zero));
  int n = 0;
  for (int w = 0; w < fa100::kTraceWarps; ++w)
    for (unsigned int i = 0; i < cnt[w] && n < max_events; ++i, ++n)
      for (int j = 0; j < 3; ++j) out[n * 3 + j] = host_buf[(w * fa100::kTracePerWarp + i) * 3 + j];
  return n;
#else
  (void)out; (void)max_events;
  return 0;
#endif
}

extern "C" int fa100_debug_prof(uint32_t* out, int max_words) {
  // per-warp cycle buckets of the timeline build (20 warps x 8 buckets); returns the number of words (0 in the product build)
#ifdef FA100_TRACE
  const int n = fa100::kTraceWarps * 8 < max_words ? fa100::kTraceWarps * 8 : max_words;
  if (cudaMemcpyFromSymbol(out, fa100::g_fa100_prof, n * sizeof(uint32_t)) != cudaSuccess) return -3;
  return n;
#else
  (void)out; (void)max_words;
  return 0;
#endif
}
