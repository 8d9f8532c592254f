/* fa100.h -- C ABI of the B200-native FlashAttention forward (libfa100.so).
 *
 * This is the drop-in boundary for the reference's forward path.  In the reference the same
 * boundary is in-process C++ (all paths relative to the reference tree):
 *   run_mha_fwd(Flash_fwd_params&, cudaStream_t)                    hopper/flash_api.cpp:367
 *   run_mha_fwd_combine(Flash_fwd_params&, cudaStream_t, bool)      hopper/flash_api.cpp:387
 *   prepare_varlen_num_blocks(params, stream, ...)                  hopper/flash.h:220
 *   get_num_splits / get_pack_gqa heuristics                        hopper/flash_api.cpp:426-470
 * and its torch-free precedent is
 *   int flash::flash_attention_forward(FlashAttentionParams&, cudaStream_t)
 *                                                                    standalone/include/flash_api.h:222-247
 * The struct below is the subset of Flash_fwd_params (hopper/flash.h:12-168) this path needs.
 *
 * Rules: plain pointers and sizes only (no torch types); the library allocates nothing, keeps no
 * pointer after a call returns, launches on the caller's stream and never synchronises, so every
 * call is re-entrant and CUDA-graph capturable.  Every function returns 0 on success or a negative
 * FA100_ERR_* code; fa100_last_error() gives the thread-local message.  Nothing ever calls exit().
 */
#ifndef FA100_H_
#define FA100_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FA100_ABI_VERSION 2  /* 2: Fa100FwdParams grew the qv fields at its end */

enum {
  FA100_OK = 0,
  FA100_ERR_INVALID = -1,     /* bad argument (shape, stride, alignment, null pointer) */
  FA100_ERR_UNSUPPORTED = -2, /* valid request this build has no kernel for (raise NotImplementedError) */
  FA100_ERR_CUDA = -3         /* CUDA driver/runtime error, text in fa100_last_error() */
};

enum { FA100_BF16 = 0, FA100_FP16 = 1, FA100_E4M3 = 2, FA100_FP32 = 3 };

typedef void* fa100_stream_t; /* cudaStream_t */

/* All strides are in ELEMENTS of the tensor's dtype; the last (head-dim) stride is always 1.
 * Layouts: q (b, sq, h, d) or packed (total_q, h, d) with cu_seqlens_q;
 * k/v (b_k, sk, h_k, d|dv), packed (total_k, h_k, d), or paged (num_pages, page_size, h_k, d);
 * o like q with dv; lse fp32 (b, h, sq) or (h, total_q). */
#define FA100_VARIANT_NO_TMA_STORE 16
#define FA100_VARIANT_FOUR_ROLE 32
#define FA100_VARIANT_SPLIT_ROWS 64

typedef struct Fa100FwdParams {
  /* tensors */
  const void* q;
  const void* k;
  const void* v;
  void* o;             /* out dtype: same as q for bf16/fp16, bf16 for e4m3 */
  float* lse;          /* (b, h, sq) contiguous, or (h, total_q) when cu_seqlens_q is set */
  int64_t q_batch_stride, q_row_stride, q_head_stride;
  int64_t k_batch_stride, k_row_stride, k_head_stride; /* batch stride == page stride when paged */
  int64_t v_batch_stride, v_row_stride, v_head_stride;
  int64_t o_batch_stride, o_row_stride, o_head_stride;

  /* split-KV partials, required when num_splits > 1 (hopper/flash_api.cpp:1094-1113):
   * o_accum fp32 (splits, b, h, sq, dv) / (splits, h, total_q, dv); lse_accum fp32 (splits, b, h, sq) / (splits, h, total_q).
   * When non-null the kernel writes ONLY these fp32 partials (each normalised by its own row sum, with its LSE) and
   * leaves o / lse untouched -- also with num_splits == 1, which is how one hop of ring attention gets a partial
   * that fa100_combine can merge.  Pass null for a plain single-pass forward. */
  float* o_accum;
  float* lse_accum;

  /* sizes */
  int32_t b;           /* batch of q */
  int32_t b_k;         /* batch of the k/v tensor (cache batch; num_pages when paged) */
  int32_t seqlen_q;    /* max seqlen_q when varlen */
  int32_t seqlen_k;    /* max seqlen_k (cache capacity for a kv-cache; page_size*max_pages when paged) */
  int32_t total_q;     /* rows of q when cu_seqlens_q is set, else 0 */
  int32_t total_k;     /* rows of k when cu_seqlens_k is set, else 0 */
  int32_t h, h_k;
  int32_t d, dv;
  int32_t dtype;       /* FA100_BF16 / FA100_FP16 / FA100_E4M3 */

  /* variable length (all int32 device pointers, may be null) */
  const int32_t* cu_seqlens_q; /* (b+1) */
  const int32_t* cu_seqlens_k; /* (b+1) */
  const int32_t* seqused_q;    /* (b)   */
  const int32_t* seqused_k;    /* (b)   cache_seqlens */
  const int32_t* leftpad_k;    /* (b)   */
  const int32_t* kv_batch_idx; /* (b)   cache_batch_idx */

  /* paged KV */
  const int32_t* page_table;   /* (b, max_pages_per_seq) */
  int64_t page_table_batch_stride;
  int32_t page_size;
  int32_t num_pages;

  /* append-KV + rotary (hopper/flash_api.cpp:1036-1092) */
  const void* k_new;           /* (b, s_new, h_k, d) or (total_k_new, h_k, d) */
  const void* v_new;
  int64_t knew_batch_stride, knew_row_stride, knew_head_stride;
  int64_t vnew_batch_stride, vnew_row_stride, vnew_head_stride;
  int32_t seqlen_knew;
  int32_t total_knew;
  const int32_t* cu_seqlens_knew;
  const void* rotary_cos;      /* (seqlen_ro, rotary_dim/2) in q dtype */
  const void* rotary_sin;
  const int32_t* seqlens_rotary;
  int32_t rotary_dim;
  int32_t is_rotary_interleaved;

  /* FP8 descale, fp32 (b, h_k) with element strides (hopper/flash_api.cpp:1115-1146); may be null */
  const float* q_descale;
  const float* k_descale;
  const float* v_descale;
  int64_t q_descale_batch_stride, q_descale_head_stride;
  int64_t k_descale_batch_stride, k_descale_head_stride;
  int64_t v_descale_batch_stride, v_descale_head_stride;

  /* math */
  float softmax_scale;
  float softcap;               /* 0 = off */
  int32_t is_causal;
  int32_t window_left;         /* -1 = unbounded */
  int32_t window_right;        /* -1 = unbounded */
  int32_t attention_chunk;     /* 0 = off */

  /* scheduling */
  int32_t num_splits;          /* >= 1 (resolved by the caller, e.g. through fa100_plan) */
  int32_t pack_gqa;            /* 0/1 (resolved by the caller) */
  int32_t num_sm;              /* SMs this call may use (multiProcessorCount - sm_margin); 0 = all */
  /* Optional device-built schedule for ragged batches: a buffer of fa100_scheduler_metadata_words(b) int32 that
   * fa100_prepare_scheduler has filled for the SAME sizes, lengths, num_splits and pack_gqa.  Null = static schedule. */
  const int32_t* scheduler_metadata;
  /* Kernel variant.  The 2-CTA kernel runs CTA pairs with cta_group::2 MMAs (M = 256 over two SMs, each CTA holding half
   * of every K / V tile); it is eligible for bf16 / fp16, d = dv = 128, K/V not paged, not the dual-stream decode kernel.
   * 0 = the library's default (today: one CTA per work item), 1 = pairs wherever eligible, 2 = one CTA per work item
   * always (low two bits).  fa100_prepare_scheduler must see the same value as fa100_fwd.
   * Flag bits for A/B measurements: FA100_VARIANT_NO_TMA_STORE = O leaves through register stores everywhere,
   * FA100_VARIANT_FOUR_ROLE = the four-role prefill mainloop of csrc/fa100_fwd2.cuh (one MMA warp issuing PV_t, QK_t in order
   * with P_t overlaid on S_t, separate softmax and correction warpgroups, P published in two pieces) instead of the default
   * (score buffer shared by the two Q tiles, QK and PV issuing warps, softmax warps own rescale + epilogue).  Parity-green on
   * the whole suite; measured 12 % slower on C2, hence opt-in.  FA100_VARIANT_SPLIT_ROWS = two softmax threads per
   * query row (csrc/fa100_fwd_split.cuh; 16 softmax warps, each thread on 64 of a block's 128 key columns): parity-green, 5-8 % slower
   * on C2, opt-in. */
  int32_t variant;
  /* Fused merge of a running partial (ring / context-parallel attention): with o_accum / lse_accum set and
   * num_splits == 1, 1 = read the fp32 partial already in o_accum / lse_accum, merge this call's result into it with the
   * combine arithmetic (LSE' = logaddexp, O' = e^(LSE-LSE') O + e^(LSE_t-LSE') O_t) and write it back in place;
   * 2 = the same, but write the merged result to o / lse (output dtype) instead.  0 = plain partial. */
  int32_t merge_partial;
  /* 1 = fa100_prepare_scheduler was called for THIS call with lse_accum set (it pre-fills the LSE partials with -inf, the
   * "empty partial" marker, in the same launch that builds the schedule): fa100_fwd then skips its own fill launch.
   * Leave 0 with a schedule built ahead of time (get_scheduler_metadata). */
  int32_t lse_accum_prefilled;
  int32_t reserved0;   /* must be 0 */
  /* Dynamic tile scheduling (role of the reference's tile_count_semaphore, hopper/tile_scheduler.hpp:363-470,
   * flash_api.cpp:996-1025): one int32 on the device that the CALLER HAS SET TO ZERO in stream order before this call and does
   * not share with another launch in flight.  With it, under causal / local masks (row blocks of unequal length) the CTAs take work
   * items from one length-ordered list as they finish instead of in a fixed round-robin; results are bitwise the same either way.
   * Null = static schedule.  Carved out of the reserved words: the struct layout and ABI version are unchanged. */
  int32_t* tile_counter;
  int32_t reserved[2]; /* must be 0 */
  /* ABI version 2.  qv (hopper/flash_api.cpp:1028-1050, params.qv_ptr): a second query tensor (b, sq, h, dv) or packed
   * (total_q, h, dv), same dtype and layout rules as q; the scores become (q k^T + qv v^T) * softmax_scale.  Like the
   * reference: 16-bit dtypes, d <= 64, dv >= 256 (this library: dv == 256).  Null = off. */
  const void* qv;
  int64_t qv_batch_stride, qv_row_stride, qv_head_stride;
} Fa100FwdParams;

/* Backward pass (hopper/flash_api.cpp:1267-1567 mha_bwd -> run_mha_bwd).  First version, scope:
 * bf16 / fp16, d == dv <= 128 (multiple of 8), MHA / GQA / MQA, fixed-length or variable-length (cu_seqlens / seqused) batches,
 * causal (bottom-right aligned), local windows, softcap; larger head dims return FA100_ERR_UNSUPPORTED.  Deterministic:
 * no atomics (dQ and dK/dV come from two kernels that each recompute the scores).  All strides in elements. */
typedef struct Fa100BwdParams {
  const void* dout; const void* q; const void* k; const void* v; const void* out;
  const float* lse;      /* (b, h, seqlen_q) fp32, natural-log units, as the forward returns it */
  void* dq; void* dk; void* dv;
  float* softmax_d;      /* workspace / output (b, h, seqlen_q_rounded) fp32: rowsum(dout * out); seqlen_q rounded up to 128 */
  int64_t q_batch_stride, q_row_stride, q_head_stride;
  int64_t k_batch_stride, k_row_stride, k_head_stride;
  int64_t v_batch_stride, v_row_stride, v_head_stride;
  int64_t o_batch_stride, o_row_stride, o_head_stride;
  int64_t do_batch_stride, do_row_stride, do_head_stride;
  int64_t dq_batch_stride, dq_row_stride, dq_head_stride;
  int64_t dk_batch_stride, dk_row_stride, dk_head_stride;
  int64_t dv_batch_stride, dv_row_stride, dv_head_stride;
  int32_t b, seqlen_q, seqlen_k, h, h_k, d, dv_dim, dtype;
  int32_t is_causal, window_left, window_right;
  float softmax_scale, softcap;
  int32_t deterministic, num_sm;
  /* variable-length batches (packed (total, h, d) tensors, batch strides unused): cu_seqlens (b + 1) and / or seqused (b);
   * seqlen_q / seqlen_k are then the maxima; lse is (h, total_q) and softmax_d (h, total_q rounded up to 128) */
  const int32_t* cu_seqlens_q; const int32_t* cu_seqlens_k; const int32_t* seqused_q; const int32_t* seqused_k;
  int32_t total_q, total_k;
  int32_t reserved[6];
} Fa100BwdParams;

/* LSE-weighted merge of split partials (hopper/flash_fwd_combine_kernel.h:200-483,
 * mha_combine hopper/flash_api.cpp:1569-1670).  Element strides; d stride is 1. */
/* `o` / `lse` may alias split 0 of `o_partial` / `lse_partial` (same strides): every row is read completely before it is
 * written, which is how ring attention keeps a running fp32 merge in place. */
typedef struct Fa100CombineParams {
  const float* o_partial;   /* (splits, b, s, h, d) logical */
  const float* lse_partial; /* (splits, b, s, h)   logical */
  void* o;                  /* (b, s, h, d) in out_dtype */
  float* lse;               /* (b, s, h) logical */
  int64_t op_split_stride, op_batch_stride, op_row_stride, op_head_stride;
  int64_t lp_split_stride, lp_batch_stride, lp_row_stride, lp_head_stride;
  int64_t o_batch_stride, o_row_stride, o_head_stride;
  int64_t lse_batch_stride, lse_row_stride, lse_head_stride;
  int32_t num_splits, b, s, h, d;
  int32_t out_dtype;        /* FA100_BF16 / FA100_FP16 / FA100_FP32 */
  /* per-batch row counts for packed (varlen) layouts; null = all s rows valid */
  const int32_t* cu_seqlens;
  const int32_t* seqused;
  int32_t reserved[8];
} Fa100CombineParams;

/* Sizes-only planning (the role of get_num_splits / get_pack_gqa / tile_size_fwd_*,
 * hopper/flash_api.cpp:415-470, hopper/heuristics.h:9-59).  Needs no device buffer. */
typedef struct Fa100Shape {
  int32_t b, seqlen_q, seqlen_k, h, h_k, d, dv, dtype;
  int32_t is_causal, is_local, is_varlen, is_paged, page_size, has_new_kv;
  int32_t num_splits;   /* 0 = choose; >= 1 = forced */
  int32_t pack_gqa;     /* -1 = choose; 0/1 = forced */
  int32_t num_sm;       /* 0 = query the current device */
} Fa100Shape;

typedef struct Fa100Plan {
  int32_t block_m;      /* query rows per CTA tile */
  int32_t block_n;      /* keys per main-loop step */
  int32_t num_splits;
  int32_t pack_gqa;
  int32_t num_sm;
  int32_t num_work;     /* CTA work items the forward kernel will iterate */
  int32_t smem_bytes;
  int32_t dual_stream;  /* 1: decode kernel whose two tile pipelines split the key blocks; num_splits (even) counts partials */
  int32_t reserved[8];
} Fa100Plan;

int fa100_abi_version(void);
const char* fa100_last_error(void);
int fa100_plan(const Fa100Shape* shape, Fa100Plan* out);
int fa100_fwd(const Fa100FwdParams* p, fa100_stream_t stream);
int fa100_combine(const Fa100CombineParams* p, fa100_stream_t stream);
/* replaces run_mha_bwd (hopper/flash_api.cpp:1267-1567, flash_bwd_launch_template.h): three launches, preprocess (softmax_d),
 * dQ, dK/dV; allocates nothing (softmax_d is the caller's). */
int fa100_bwd(const Fa100BwdParams* p, fa100_stream_t stream);

/* Split-KV forward + combine with NO memory allocation anywhere: the fp32 partials (the only scratch memory the forward
 * uses) live in a static device array inside the library instead of caller-supplied o_accum / lse_accum.
 * No counterpart in the reference: added for callers that cannot allocate device
 * memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 *   - p as for fa100_fwd, with o and lse set, o_accum = lse_accum = NULL, num_splits in [2, 256], dense (not packed) q;
 *     result = fa100_fwd with partials followed by fa100_combine into o (bf16 / fp16) and lse (b, h, seqlen_q).
 *   - limit: num_splits * b * h * seqlen_q * (dv + 1) * 4 bytes <= FA100_NOALLOC_SCRATCH_BYTES (16 MiB); beyond it the call
 *     returns FA100_ERR_INVALID with a message and launches nothing.  Every GPU that loads the library holds the array
 *     until the process exits.
 *   - all calls share the array and it keeps the previous call's contents (the LSE slots are reset at the start of each
 *     call); calls that overlap on the GPU (several streams, concurrent graph replays) RACE on it: serialise them.
 *   - calls no cudaMalloc / cudaFree / allocating routine; stream-ordered, no host synchronisation. */
#define FA100_NOALLOC_SCRATCH_BYTES (16 * 1024 * 1024)
int fa100_fwd_splitkv_noalloc(const Fa100FwdParams* p, fa100_stream_t stream);
int64_t fa100_noalloc_scratch_bytes(void);

/* Device-side schedule for ragged batches: the role of prepare_varlen_num_blocks (hopper/flash.h:220,
 * hopper/flash_prepare_scheduler.cu:43-215) and of the metadata tensor mha_fwd_get_scheduler_metadata returns
 * (hopper/flash_api.cpp:521-662, layout :996-1025).  One small kernel on `stream` reads cu_seqlens / seqused / leftpad
 * of `p` (q, k, v, o may be null), counts every batch element's row blocks and key blocks, orders the batch elements
 * longest first, gives each its own split count (<= num_splits, sized so that every split holds about the same number of
 * key blocks) and writes, after an 8-word header the caller owns, [work items, b, order[b], prefix[b + 1], nsplit[b], scratch].
 * fa100_fwd then walks that list instead of the static (row block, batch, head) grid: no entries for rows past a
 * sequence's end, expensive sequences first.  `metadata` must hold fa100_scheduler_metadata_words(p->b) int32. */
int64_t fa100_scheduler_metadata_words(int32_t b);
int fa100_prepare_scheduler(const Fa100FwdParams* p, int32_t* metadata, int64_t metadata_words, fa100_stream_t stream);
/* Append-KV + rotary pass that precedes fa100_fwd when flash_attn_with_kvcache is given k / v (and rotary_cos / sin).
 * Replaces the AppendKV phases of the reference kernel: store_kv_new and the Q rotation
 * (hopper/mainloop_fwd_sm90_tma_gmma_ws.hpp:1447-1713, :1098-1123; hopper/rotary.h:17-485; positions
 * hopper/seqlen.h:60-93).  Reads p->k_new / v_new (+ strides, seqlen_knew / cu_seqlens_knew), p->seqused_k (lengths
 * BEFORE the append), p->kv_batch_idx / page_table, p->rotary_* and p->seqlens_rotary, p->q (+ strides, cu_seqlens_q,
 * seqused_q), p->is_causal / window_* / attention_chunk (causal or local: query row i is rotated at position base + i,
 * else every query row at base).  Writes the new rows into p->k / p->v in place, the rotated copy of q into `q_rotated`
 * (element strides given; ignored without rotary; the caller's q is never modified) and seqused_k + n_new into
 * `seqused_k_out` (b int32).  The caller then runs fa100_fwd with q = q_rotated and seqused_k = seqused_k_out. */
int fa100_append_kv(const Fa100FwdParams* p, void* q_rotated, int64_t qr_batch_stride, int64_t qr_row_stride,
                    int64_t qr_head_stride, int32_t* seqused_k_out, fa100_stream_t stream);
/* self-test hook used by the GPU tests: returns the hang-watchdog record of the last trapped kernel */
int fa100_debug_hang_info(uint32_t out[4]);
/* profiling hook: events {tag | warp << 16, begin_clk, end_clk} of the opt-in timeline build (libfa100_trace.so,
 * -DFA100_TRACE); returns the event count, always 0 from the product library */
int fa100_debug_trace(uint32_t* out, int max_events);
/* profiling hook: per-warp cycle buckets (8 per warp) of the timeline build; returns words written, 0 from the product library */
int fa100_debug_prof(uint32_t* out, int max_words);

#ifdef __cplusplus
}
#endif
#endif /* FA100_H_ */
