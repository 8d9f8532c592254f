// fa100_fwd_kernel.cuh -- warp-specialised FlashAttention forward for sm_100a.
//
// Role of the reference's mainloop + softmax + mask + epilogue + tile scheduler
// (hopper/mainloop_fwd_sm90_tma_gmma_ws.hpp:591-1352, softmax.h:92-168, mask.h:47-162, block.h:14-60,
// epilogue_fwd.hpp:215-481, tile_scheduler.hpp), re-designed for Blackwell:
//
//   warps 0-3   softmax for Q tile 0: one thread == one query row == one TMEM lane (row max / sum without shuffles)
//   warps 4-7   softmax for Q tile 1
//   warp  8     QK issuer (elected lane): S = Q_t K^T (SS form) into the score buffer the two tiles share
//   warp  9     TMA producer (elected lane): Q tiles once per work item, K/V tiles through an mbarrier ring
//   warp  10    PV issuer (elected lane): O_t += P_t V (TS form, P in TMEM)
//   warp  11    geometry: decodes the work items ahead of everybody else into a shared-memory ring (and completes the third
//               warpgroup for setmaxnreg: softmax warps run with 224 registers, the rest 56)
//
// TMEM (512 columns): S @0 (fp32 scores; ONE buffer the two Q tiles take turns in), P_0 @128, P_1 @192 (operand dtype: 64
// columns at 16 bit, 32 at e4m3), O_0 @256, O_1 @384.  Two 128-row Q tiles per CTA ping-pong through the tensor pipe: a
// softmax warpgroup copies its scores into registers and hands S to the other tile at once, so the next S_t never waits for
// P_t V.  O is only rescaled when the running row max grew by more than 2^8 (exact: the final 1/rowsum and LSE use the
// same stale max); the rescale is done in place in TMEM by the softmax warps, which also run the epilogue (normalise,
// convert, store O and LSE).  The opt-in four-role mainloop with a separate correction warpgroup is csrc/fa100_fwd2.cuh.
#pragma once
#include "fa100_ptx.cuh"

namespace fa100 {

constexpr int kBlockM = 128;   // query rows per Q tile
constexpr int kBlockN = 128;   // keys per main-loop step
constexpr int kQTiles = 2;     // Q tiles per CTA
constexpr int kSoftmaxWarps = 8;
constexpr int kMmaWarp = 8;
constexpr int kLoadWarp = 9;
constexpr int kPvWarp = 10;
constexpr int kGeomWarp = 11;  // computes every work item's geometry ahead of the other roles (geom ring)
constexpr int kThreads = 12 * 32;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kRescaleThreshold = 8.0f;  // log2 units
// e4m3 P = 2^(x - m + kFp8Offset): the offset keeps small probabilities out of the subnormal range (hopper/softmax.h:18-40
// uses 8 with an exact running max).  Here the max may lag by kFp8RescaleThreshold, so the largest P is
// 2^(offset + threshold) = 2^8.75 = 430 < 448 with offset 6, threshold 2.75 (e4m3 max; the cvt saturates anyway).  A threshold of 0 made every early key
// block rescale O in tensor memory (any row of the warp with a new max): 325 clk per block on random data.
#ifndef FA100_FP8_OFFSET
#define FA100_FP8_OFFSET 6.0f
#define FA100_FP8_THRESHOLD 2.75f
#endif
constexpr float kFp8Offset = FA100_FP8_OFFSET, kFp8RescaleThreshold = FA100_FP8_THRESHOLD;

struct FwdKernelParams {
  void* o;
  float* lse;
  float* o_accum;
  float* lse_accum;
  int64_t o_batch_stride, o_row_stride, o_head_stride;            // elements
  int64_t oacc_split_stride, oacc_batch_stride, oacc_head_stride; // fp32 elements; row stride == dv
  int64_t lse_batch_stride, lse_head_stride;                      // fp32 elements; row stride == 1
  int64_t lseacc_split_stride;
  const int32_t* cu_seqlens_q;
  const int32_t* cu_seqlens_k;
  const int32_t* seqused_q;
  const int32_t* seqused_k;
  const int32_t* leftpad_k;
  const int32_t* kv_batch_idx;
  const int32_t* page_table;
  int64_t page_table_batch_stride;
  const float* q_descale;
  const float* k_descale;
  const float* v_descale;
  int64_t q_descale_batch_stride, q_descale_head_stride;
  int64_t k_descale_batch_stride, k_descale_head_stride;
  int64_t v_descale_batch_stride, v_descale_head_stride;
  int32_t b, h, h_k, seqlen_q, seqlen_k, d, dv;
  int32_t page_size;      // 0 = not paged
  int32_t kv_box_rows;    // rows per K/V TMA box: 128, or gcd(128, page_size) when paged
  int32_t qk_boxes;       // 128-byte column boxes of Q / K actually loaded: ceil(d * elt / 128) <= HD * elt / 128
  int32_t v_boxes;        // same for V: ceil(dv * elt / 128)
  int32_t qk_slices;      // MMA k-slices of S = Q K^T actually issued: ceil(d / (32 / elt))
  int32_t pv_n;           // N of O += P V: dv rounded up to 16
  int32_t qv_slices;      // qv (hdim-256 instance only): MMA k-slices of the extra score term Qv V^T = ceil(dv / 16); 0 = off
  int32_t qheads_per_kv;  // G
  int32_t pack;           // Gp: G when pack_gqa else 1
  int32_t pos_per_tile;   // 128 / Gp query positions per Q tile
  int32_t num_m_blocks;   // ceil(max_seqlen_q / (2 * pos_per_tile))
  int32_t heads_work;     // h_k when packed else h
  int32_t num_splits;
  int32_t num_work;       // static schedule: every (m block, batch, head, split); an upper bound when sched is set
  int32_t tiles_per_item; // 128-row Q tiles one work item covers: 2, or 1 in the dual-stream and head-dim-256 kernels
  int32_t lpt_groups;     // > 0 (causal / local with a tile counter): (batch, KV head) groups per super-group of the length-ordered work list
  // Device-built schedule for ragged batches (fa100_prepare_scheduler): [0] work items, [1] batches, [2 .. 2+b) batch
  // ids, most key blocks first, [2+b .. 2+2b] exclusive prefix of work items over that order, [3+2b .. 3+3b) CTA-level
  // split count of each batch element in that order (<= num_splits).  Null: static schedule.
  const int32_t* sched;
  int32_t* tile_counter;  // dynamic scheduling (zeroed by the caller): the geometry warp takes the CTA's next item with an atomicAdd; null = static
  int32_t num_sm;         // SMs the launch may use (the schedule kernel sizes the per-batch split counts with it)
  int32_t is_causal, is_local, window_left, window_right, attention_chunk;
  int32_t out_fp16;       // 1: fp16 output, 0: bf16 output
  int32_t o_tma;          // 1: O leaves through shared memory and TMA stores (no pack-GQA, no partials)
  int32_t o_tma_check;    // 1 (packed or seqused_q batches): only for tiles that lie wholly inside their sequence
  int32_t merge;          // 1 / 2: merge with the fp32 partial already in o_accum / lse_accum, in place / into o and lse
  float scale_log2;       // softmax_scale * log2(e)   (softcap: softcap * log2(e))
  float softcap_pre;      // softmax_scale / softcap, 0 when softcap is off
  float softmax_scale;
};

// ----------------------------------------------------------------------------- geometry shared by all roles
struct WorkItem { int m_blk, head, bi, split, nsplits; };  // nsplits: CTA-level splits of this batch element

struct SeqInfo {
  int sq, sk;      // this batch element's lengths
  int q_off;       // row offset of the sequence inside a packed q / o / lse
  int k_off;       // row offset inside k/v (packed offset + left pad)
  int bq, bk;      // batch coordinates for TMA (0 when packed)
};

struct TileRange { int n_min, n_max; bool has_rows; };

// Work-item geometry every role needs (make_geom in fa100_fwd_roles.cuh).  In the default kernel the otherwise idle twelfth
// warp computes it one to four items ahead and hands it to the other roles through shared memory (geom ring below): a
// make_geom costs ~2500-3000 clk of integer divisions and dependent loads, twice per item on every softmax warp before
// (cycle-bucket build with -DFA100_PROF_EPI: 6500 of the 36800 clk a softmax warp spent per 1024-token causal item).
struct WorkGeom {
  WorkItem it;
  SeqInfo seq;
  TileRange tr[2];
  int n_lo, n_hi;  // union of the tiles' key-block ranges; n_hi <= n_lo means nothing to do
  int hk;          // kv head
  int w;           // index of the item in the work list; -1 in the geometry ring = the list is exhausted
};
constexpr int kGeomSlots = 4;

// Static persistent schedule: CTA c runs list positions c, 2 grid - 1 - c, 2 grid + c, ... (boustrophedon).  Item lengths
// fall linearly inside a K/V group, so pairing a forward pass with a reversed one evens out what plain round-robin
// leaves (makespan / mean 1.11 -> 1.05 at C3, measured on the host model in DESIGN.md section 4.1).
// PAIR: the two CTAs of a cluster walk the list together (one entry = four 128-row Q tiles, two per CTA).
template <bool PAIR = false>
FA_DEVICE int work_index(int round) {
  const int grid = PAIR ? int(gridDim.x) >> 1 : int(gridDim.x), cta = PAIR ? int(blockIdx.x) >> 1 : int(blockIdx.x);
  return round * grid + ((round & 1) ? grid - 1 - cta : cta);
}

// Header in front of the device-built schedule: [magic, num_splits, pack_gqa, num_sm, num_work bound] from the host
// (get_scheduler_metadata; informational) and [magic, .., .., num_sm, .., tiles_per_item, CTA-level splits, pack] from
// fa100_schedule_kernel: the forward kernel refuses a schedule built for another plan (sched_matches).
constexpr int kSchedHeaderWords = 8;
constexpr int32_t kSchedMagic = 0xFA100;
FA_DEVICE bool sched_matches(const FwdKernelParams& p) {
  const int32_t* h = p.sched - kSchedHeaderWords;
  return h[0] == kSchedMagic && h[3] == p.num_sm && h[5] == p.tiles_per_item && h[6] == p.num_splits && h[7] == p.pack && p.sched[1] == p.b;
}
FA_DEVICE int total_work(const FwdKernelParams& p) { return p.sched ? p.sched[0] : p.num_work; }

FA_DEVICE WorkItem decode_work(const FwdKernelParams& p, int w) {
  // w = ((((bi * h_k + hk) * m_blocks + m_rev) * heads_per_group + g) * splits + split.  Everything that reads one
  // (batch, KV head) -- its splits, the query heads of the group and all their row blocks -- is adjacent, so the ~148
  // work items in flight share one or two K/V heads (4 MB at C2) in L2 instead of streaming all of K/V at once
  // (the reference swizzles heads by L2 size for the same reason, hopper/tile_scheduler.hpp:305-360).  Inside a group
  // the long (late) causal row blocks come first.
  WorkItem it;
  const int gsz = p.heads_work / p.h_k;  // query heads per KV head that are separate work items (1 when packed)
  if (p.sched) {
    // ragged batch: batches in the order the schedule kernel chose (most key blocks first), each with its own number
    // of row blocks and no entries for rows past its end; inside a batch the order below is unchanged
    const int32_t* order = p.sched + 2;
    const int32_t* prefix = order + p.b;
    int lo = 0, hi = p.b - 1;  // largest r with prefix[r] <= w
    while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (prefix[mid] <= w) lo = mid; else hi = mid - 1; }
    const int per_batch = prefix[lo + 1] - prefix[lo];
    it.nsplits = min(prefix[p.b + 1 + lo], p.num_splits);  // never past the partials the caller allocated
    const int m_blocks = per_batch / (p.heads_work * it.nsplits);
    w -= prefix[lo];
    it.bi = order[lo];
    it.split = w % it.nsplits;  w /= it.nsplits;
    const int g = w % gsz;        w /= gsz;
    if (p.tile_counter) {
      // dynamic schedule: a batch element's row blocks longest first ACROSS its heads (taken in list order, a long block at the end
      // of the list would be the tail: host model 1.076 -> 1.02 for 16 x 16 heads x 2048 causal)
      const int groups = p.heads_work / gsz;
      it.head = (w % groups) * gsz + g;
      it.m_blk = m_blocks - 1 - w / groups;
      return it;
    }
    it.m_blk = m_blocks - 1 - w % m_blocks;  w /= m_blocks;
    it.head = w * gsz + g;
    return it;
  }
  it.nsplits = p.num_splits;
  it.split = w % p.num_splits;  w /= p.num_splits;
  const int g = w % gsz;        w /= gsz;
  if (p.lpt_groups > 0) {
    // Causal / local masks make the row blocks of a sequence unequal (1 .. num_m_blocks key-block pairs), and the order above
    // hands a CTA whole stretches of one group: makespan / mean 1.045 at C2, 1.11 for 4 x 16 heads x 8192 (host model,
    // DESIGN.md 4.1).  Here the list is cut into super-groups of lpt_groups (batch, KV head) groups -- about one round of the
    // grid -- and inside a super-group the row blocks go longest first ACROSS its groups, so the items of one round have
    // (nearly) the same length and every CTA's sum comes out the same.  The query heads of a KV head stay adjacent (they
    // run in the same round and share its K / V in L2).
    const int groups = p.b * p.h_k;
    const int per_sg = p.num_m_blocks * p.lpt_groups;
    const int sg = w / per_sg, rem = w - sg * per_sg;
    const int in_sg = min(p.lpt_groups, groups - sg * p.lpt_groups);  // the last super-group may be smaller
    const int m_rev = rem / in_sg, group = sg * p.lpt_groups + (rem - m_rev * in_sg);
    it.m_blk = p.num_m_blocks - 1 - m_rev;
    it.bi = group / p.h_k;
    it.head = (group - it.bi * p.h_k) * gsz + g;
    return it;
  }
  it.m_blk = p.num_m_blocks - 1 - w % p.num_m_blocks;  w /= p.num_m_blocks;
  it.head = (w % p.h_k) * gsz + g;
  it.bi = w / p.h_k;
  return it;
}

FA_DEVICE SeqInfo load_seq(const FwdKernelParams& p, int bi) {
  SeqInfo s;
  const int leftpad = p.leftpad_k ? p.leftpad_k[bi] : 0;
  s.q_off = p.cu_seqlens_q ? p.cu_seqlens_q[bi] : 0;
  s.k_off = (p.cu_seqlens_k ? p.cu_seqlens_k[bi] : 0) + leftpad;
  s.sq = p.seqused_q ? p.seqused_q[bi] : (p.cu_seqlens_q ? p.cu_seqlens_q[bi + 1] - p.cu_seqlens_q[bi] : p.seqlen_q);
  s.sk = (p.seqused_k ? p.seqused_k[bi] : (p.cu_seqlens_k ? p.cu_seqlens_k[bi + 1] - p.cu_seqlens_k[bi] : p.seqlen_k)) - leftpad;
  s.bq = p.cu_seqlens_q ? 0 : bi;
  s.bk = p.cu_seqlens_k ? 0 : (p.kv_batch_idx ? p.kv_batch_idx[bi] : bi);
  return s;
}

FA_DEVICE int floor_mod(int x, int m) { int r = x % m; return r < 0 ? r + m : r; }

// Key-block range of one Q tile (arithmetic of hopper/block.h:14-60, on 128-row tiles; -1 windows = unbounded).
FA_DEVICE TileRange tile_range(const FwdKernelParams& p, const SeqInfo& s, int pos0, int split, int nsplits) {
  TileRange r;
  r.has_rows = pos0 < s.sq;
  r.n_min = 0; r.n_max = 0;
  if (!r.has_rows) return r;
  const int pos_end = min(pos0 + p.pos_per_tile, s.sq);
  const int shift = s.sk - s.sq;
  int n_max = (s.sk + kBlockN - 1) / kBlockN;
  int n_min = 0;
  if (p.is_causal || p.is_local) {
    if (p.is_causal || p.window_right >= 0 || p.attention_chunk > 0) {
      int right = pos_end + shift + ((p.is_local && p.window_right >= 0) ? p.window_right : 0);
      if (p.is_local && p.window_right < 0) right = s.sk;
      if (p.is_local && p.attention_chunk > 0) {
        const int x = pos_end - 1 + shift;  // last row's absolute position
        right = min(right, x - floor_mod(x, p.attention_chunk) + p.attention_chunk);
      }
      n_max = min(n_max, right <= 0 ? 0 : (right + kBlockN - 1) / kBlockN);
    }
    if (p.is_local) {
      int left = (p.window_left >= 0) ? pos0 + shift - p.window_left : 0;
      if (p.attention_chunk > 0) { const int x = pos0 + shift; left = max(left, x - floor_mod(x, p.attention_chunk)); }
      n_min = max(0, left / kBlockN);
    }
  }
  if (nsplits > 1) {
    const int per = n_max <= n_min ? 0 : (n_max - n_min + nsplits - 1) / nsplits;
    n_min = n_min + split * per;
    n_max = min(n_min + per, n_max);
  }
  r.n_min = n_min; r.n_max = max(n_max, n_min);
  return r;
}

// ----------------------------------------------------------------------------- shared memory
template <int HD, int ELT, bool PAIR = false>
struct FwdSmem {
  static constexpr int kQBytes = kBlockM * HD * ELT;   // one Q tile
  // one K or V tile; a CTA of a pair holds half of it (K: 64 of the 128 keys, V: 64 of the 128 value columns)
  static constexpr int kKVBytes = kBlockN * HD * ELT / (PAIR ? 2 : 1);
  static constexpr int kBarrierBytes = 1024;           // mbarriers + TMEM base (+ the trace build's counters in the last 128 B)
  static constexpr int kBudget = 227 * 1024 - kBarrierBytes;
  // head dims above 128 run one Q tile per work item: O alone needs 256 TMEM columns, so the second tile pipeline stays
  // idle (SOLO geometry in fa100_fwd_roles.cuh) and only one Q slot is carved out
  static constexpr int kQSlots = HD > 128 ? 1 : kQTiles;
  // O staging for the TMA-store epilogue: per Q tile one [128 rows][128 B] box (64 output columns of 16 bits), reused for
  // every 64-column slice of the tile.  The hdim-256 instance has no room for it and keeps the register stores.
  static constexpr int kOStageBytes = kBlockM * 128;
  static constexpr int kOBytes = HD > 128 ? 0 : kQTiles * kOStageBytes;
  static constexpr int kStagesRaw = ((kBudget - kQSlots * kQBytes - kOBytes) / kKVBytes) & ~1;  // even: half K stages, half V stages
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static_assert(kStages >= 2, "need at least one K and one V stage");
  static constexpr int kDataBytes = kQSlots * kQBytes + kStages * kKVBytes + kOBytes;
  static constexpr int kTotalBytes = kDataBytes + kBarrierBytes;
};

struct FwdBarriers {
  uint64_t q_full[kQTiles], q_empty[kQTiles];
  uint64_t s_full[kQTiles], p_full[kQTiles], pv_done[kQTiles];
  uint64_t s_free;
  uint64_t o_full[kQTiles], o_empty[kQTiles];
  uint64_t kv_full[8], kv_empty[8];
  uint64_t geom_full[kGeomSlots], geom_empty[kGeomSlots];  // geometry ring: filled by the geometry warp, released by every consumer warp
  WorkGeom geom[kGeomSlots];
  uint32_t tmem_base;
  uint32_t ts[8];  // latency-probe timestamps of the profiling build (0,1: QK_t issued; 2: s_free arrive; 4,5: p_full[t] arrive)
  uint32_t pad_[3];
};
static_assert(sizeof(FwdBarriers) <= 1024 - 128, "barrier block outgrew its slot");

}  // namespace fa100
