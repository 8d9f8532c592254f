// fa100_fwd_roles.cuh -- producer (TMA) and MMA-issuer roles of the forward kernel.
#pragma once
#include "fa100_fwd_kernel.cuh"

namespace fa100 {

// Work-item geometry (struct WorkGeom, fa100_fwd_kernel.cuh): every role derives it identically with make_geom, or -- RING --
// takes it from the shared-memory ring the geometry warp fills.

// DUAL (decode, all query rows of an item fit one 128-row tile): both TMEM tile pipelines work on the SAME rows and
// split the item's key blocks between them -- tile 0 takes blocks n_hi-1, n_hi-3, ..., tile 1 takes n_hi-2, n_hi-4, ... --
// and each writes its own fp32 partial (split index 2 * split + t) for the combine kernel.  Without it the second
// pipeline idles and the per-block chain QK -> softmax -> PV (~3300 clk) is longer than the 64 KB block takes to
// arrive from HBM (~2450 clk at 6.5 TB/s over 148 SMs).
// SOLO (head dim > 128): a work item is ONE 128-row tile; tile 1 never has rows.
// PAIR (cta_group::2): a work item is four 128-row tiles, tile t of CTA r being row block 4 m + 2 t + r.  The t-th tiles
// of the two CTAs form one M = 256 MMA, so both walk the UNION of their key-block ranges; a block outside a CTA's own
// range is simply fully masked for its rows (at most one block per tile pair under a causal mask).
template <bool DUAL, bool SOLO = false, bool PAIR = false>
FA_DEVICE WorkGeom make_geom(const FwdKernelParams& p, int w, int rank = 0) {
  WorkGeom g;
  g.w = w;
  g.it = decode_work(p, w);
  g.seq = load_seq(p, g.it.bi);
  g.hk = p.pack > 1 ? g.it.head : g.it.head / p.qheads_per_kv;
  g.n_lo = 0x7fffffff; g.n_hi = 0;
  if (PAIR) {
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      // the union of the two whole ranges is what gets cut into splits: both CTAs must walk the same blocks, and a block
      // may belong to one split only (cutting each tile's own range, as the 1-CTA kernel does, would give the two
      // CTAs different shares of the same split)
      const TileRange a = tile_range(p, g.seq, (g.it.m_blk * 4 + 2 * t + rank) * p.pos_per_tile, 0, 1);
      const TileRange b = tile_range(p, g.seq, (g.it.m_blk * 4 + 2 * t + (rank ^ 1)) * p.pos_per_tile, 0, 1);
      const bool la = a.n_max > a.n_min, lb = b.n_max > b.n_min;
      TileRange u;
      u.has_rows = a.has_rows || b.has_rows;
      u.n_min = la && lb ? min(a.n_min, b.n_min) : (la ? a.n_min : b.n_min);
      u.n_max = la && lb ? max(a.n_max, b.n_max) : (la ? a.n_max : b.n_max);
      if (!la && !lb) u.n_min = u.n_max = 0;
      if (g.it.nsplits > 1) {
        const int per = u.n_max <= u.n_min ? 0 : (u.n_max - u.n_min + g.it.nsplits - 1) / g.it.nsplits;
        u.n_min = u.n_min + g.it.split * per;
        u.n_max = max(min(u.n_min + per, u.n_max), u.n_min);
      }
      g.tr[t] = u;
      if (u.n_max > u.n_min) { g.n_lo = min(g.n_lo, u.n_min); g.n_hi = max(g.n_hi, u.n_max); }
    }
    return g;
  }
  if (DUAL || SOLO) {
    const TileRange r = tile_range(p, g.seq, g.it.m_blk * p.pos_per_tile, g.it.split, g.it.nsplits);
    g.tr[0] = r;
    g.tr[1] = r;
    if (SOLO) { g.tr[1].has_rows = false; g.tr[1].n_min = g.tr[1].n_max = 0; }
    else g.tr[1].n_max = max(r.n_min, r.n_max - 1);
    if (r.n_max > r.n_min) { g.n_lo = r.n_min; g.n_hi = r.n_max; }
    return g;
  }
#pragma unroll
  for (int t = 0; t < kQTiles; ++t) {
    g.tr[t] = tile_range(p, g.seq, (g.it.m_blk * kQTiles + t) * p.pos_per_tile, g.it.split, g.it.nsplits);
    if (g.tr[t].n_max > g.tr[t].n_min) { g.n_lo = min(g.n_lo, g.tr[t].n_min); g.n_hi = max(g.n_hi, g.tr[t].n_max); }
  }
  return g;
}
template <bool ONE_TILE, bool PAIR = false> FA_DEVICE int tile_pos0(const FwdKernelParams& p, const WorkItem& it, int t, int rank = 0) {
  if (PAIR) return (it.m_blk * 4 + 2 * t + rank) * p.pos_per_tile;
  return (ONE_TILE ? it.m_blk : it.m_blk * kQTiles + t) * p.pos_per_tile;  // ONE_TILE = DUAL or SOLO
}
FA_DEVICE bool tile_live(const WorkGeom& g, int t) { return g.tr[t].n_max > g.tr[t].n_min; }
FA_DEVICE bool tile_uses(const WorkGeom& g, int t, int nb) { return nb >= g.tr[t].n_min && nb < g.tr[t].n_max; }

// RING: item wi's geometry from the ring (every consumer warp takes and releases every item, in order, exactly once)
template <bool DUAL, bool SOLO, bool PAIR, bool RING, typename Bar>
FA_DEVICE WorkGeom item_geom(const FwdKernelParams& p, Bar* bar, int w, int wi, int rank) {
  if constexpr (RING) {
    const int slot = wi % kGeomSlots;
    mbar_wait(&bar->geom_full[slot], (wi / kGeomSlots) & 1, 0x90);
    return bar->geom[slot];
  } else {
    return make_geom<DUAL, SOLO, PAIR>(p, w, rank);
  }
}
template <bool RING, typename Bar>
FA_DEVICE void item_geom_release(Bar* bar, int wi) {
  if constexpr (RING) {
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(&bar->geom_empty[wi % kGeomSlots]);
  }
}
// the geometry warp: walks the CTA's work list ahead of everybody else.  Static schedule: list positions work_index(0), (1), ...;
// with FwdKernelParams.tile_counter (not for CTA pairs, whose two CTAs must agree on the item): the CTA's own index first, then
// whatever an atomicAdd on the counter hands out (hopper/tile_scheduler.hpp:363-470, DynamicPersistentTileScheduler) -- the
// round trip is hidden by the ring's depth.  A slot with w = -1 ends every consumer's loop.
template <bool DUAL, bool SOLO, bool PAIR>
FA_DEVICE void geom_role(const FwdKernelParams& p, FwdBarriers* bar, int rank) {
  const int lane = threadIdx.x & 31;
  const int n_work = total_work(p);
  const bool dynamic = !PAIR && p.tile_counter != nullptr;
  for (int wi = 0;; ++wi) {
    const int slot = wi % kGeomSlots;
    // static: as far ahead as the ring is deep.  Dynamic: ONE item ahead (item wi is claimed when every consumer has released
    // item wi - 2, i.e. while wi - 1 is being worked on) -- a CTA sitting on three claimed items at the end of the list while
    // others idle cost 10 % on 2048-item causal batches
    const int ahead = dynamic ? 2 : kGeomSlots;
    if (wi >= ahead) mbar_wait(&bar->geom_empty[(wi - ahead) % kGeomSlots], ((wi - ahead) / kGeomSlots) & 1, 0x91);
    int w = 0;
    if (lane == 0) w = dynamic ? (wi == 0 ? int(blockIdx.x) : int(gridDim.x) + atomicAdd(p.tile_counter, 1)) : work_index<PAIR>(wi);
    w = __shfl_sync(0xffffffffu, w, 0);
    const bool done = w >= n_work;
    if (lane == 0) {
      if (done) bar->geom[slot].w = -1;
      else bar->geom[slot] = make_geom<DUAL, SOLO, PAIR>(p, w, rank);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&bar->geom_full[slot]);
    if (done) break;
  }
}

struct RingState { int stage = 0; uint32_t phase = 0; };
template <int STAGES> FA_DEVICE void ring_advance(RingState& r) { if (++r.stage == STAGES) { r.stage = 0; r.phase ^= 1; } }

// ----------------------------------------------------------------------------- TMA producer
// The whole warp walks the schedule (warp-uniform control flow keeps descriptors and coordinates in uniform
// registers); one elected lane issues the bulk copies.
template <int HD, int ELT, bool DUAL, bool PAIR = false, typename Bar = FwdBarriers, bool RING = false>
FA_DEVICE void producer_role(const FwdKernelParams& p, const CUtensorMap* tmQ, const CUtensorMap* tmK,
                             const CUtensorMap* tmV, uint8_t* sQ, uint8_t* sKV, Bar* bar, int rank = 0, const CUtensorMap* tmQv = nullptr) {
  using L = FwdSmem<HD, ELT, PAIR>;
  constexpr int kRowBytes = 128;                       // 128B-swizzled rows
  constexpr int kColsPerBox = kRowBytes / ELT;         // 64 (16-bit) or 128 (8-bit)
  constexpr int kBoxesD = HD / kColsPerBox > 0 ? HD / kColsPerBox : 1;  // boxes along the head dim
  constexpr int kHalfBytes = kBlockN * kRowBytes;      // one [128 rows][128 B] region
  // K tiles cycle through the first half of the stages, V tiles through the second half: every stage (and its pair of
  // mbarriers) then has exactly one consumer warp.  With one mixed ring the kind of tile a stage holds changes from fill
  // to fill in DUAL mode, a warp that only watches "its" fills skips phases of the barrier, and the parity test aliases.
  constexpr int kRing = L::kStages / 2;
  RingState rings[2];
  uint32_t q_it[kQTiles] = {0, 0};
  // only the boxes that hold real head-dim columns are loaded (d = 192 in the 256 instance: 3 of 4)
  // qv (hdim-256 instance, d <= 64): the Q slot holds the Qv tile (dv columns) and the Q tile moves into box 1 of the K
  // stage, which K tiles of d <= 64 never touch (they fill box 0 only); see mma_role
  const bool with_qv = HD > 128 && p.qv_slices > 0;
  const uint32_t q_tx = uint32_t((p.qk_boxes + (with_qv ? p.v_boxes : 0)) * kRowBytes) * p.pack * p.pos_per_tile;
  const int n_work = total_work(p);
  for (int wi = 0;; ++wi) {
    const int w = RING ? 0 : work_index<PAIR>(wi);
    if (!RING && w >= n_work) break;
    const WorkGeom g = item_geom<DUAL, (HD > 128), PAIR, RING>(p, bar, w, wi, rank);
    item_geom_release<RING>(bar, wi);  // g is a copy
    if (RING && g.w < 0) break;
    if (g.n_hi <= g.n_lo) continue;
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      if (!tile_live(g, t)) continue;
      mbar_wait(&bar->q_empty[t], (q_it[t] & 1) ^ 1, 0x10 + t);
      const int pos0 = tile_pos0<DUAL || (HD > 128), PAIR>(p, g.it, t, rank);
      if (elect_one()) {
        // PAIR: both CTAs' Q tiles are counted on the leader's barrier; the leader announces the bytes of both
        if (!PAIR || rank == 0) mbar_arrive_expect_tx(&bar->q_full[t], PAIR ? 2 * q_tx : q_tx);
        if (with_qv) {
          tma_load_5d(sKV + kHalfBytes, tmQ, &bar->q_full[t], 0, 0, g.seq.q_off + pos0, g.it.head, g.seq.bq);
#pragma unroll
          for (int hb = 0; hb < kBoxesD; ++hb)
            if (hb < p.v_boxes) tma_load_5d(sQ + hb * kHalfBytes, tmQv, &bar->q_full[t], hb * kColsPerBox, 0, g.seq.q_off + pos0, g.it.head, g.seq.bq);
        } else
#pragma unroll
        for (int hb = 0; hb < kBoxesD; ++hb)
          if (hb < p.qk_boxes) {
            if (PAIR) tma_load_5d_2sm(sQ + t * L::kQBytes + hb * kHalfBytes, tmQ, &bar->q_full[t], hb * kColsPerBox, 0,
                                      g.seq.q_off + pos0, g.it.head, g.seq.bq);
            else tma_load_5d(sQ + t * L::kQBytes + hb * kHalfBytes, tmQ, &bar->q_full[t], hb * kColsPerBox, 0,
                             g.seq.q_off + pos0, g.it.head, g.seq.bq);
          }
      }
      __syncwarp();
      ++q_it[t];
    }
    // one K or V tile of key block nb into the next ring stage
    auto load_tile = [&](int kv, int nb) {
      const CUtensorMap* tm = kv == 0 ? tmK : tmV;
      RingState& ring = rings[kv];
      const int stage = kv * kRing + ring.stage;
      mbar_wait(&bar->kv_empty[stage], ring.phase ^ 1, 0x20 + kv);
      uint8_t* dst = sKV + stage * L::kKVBytes;
      const int boxes = kv == 0 ? p.qk_boxes : p.v_boxes;
      if (PAIR) {
        // this CTA's half of the tile: keys [64 rank, +64) of K (boxes of 64 rows), value columns [64 rank, +64) of V
        // (one box of 128 rows); the bytes of both halves are counted on the leader's barrier
        if (elect_one()) {
          if (rank == 0) mbar_arrive_expect_tx(&bar->kv_full[stage], 2 * (kv == 0 ? boxes * (kHalfBytes / 2) : kHalfBytes));
          if (kv == 0) {
#pragma unroll
            for (int hb = 0; hb < kBoxesD; ++hb)
              if (hb < boxes) tma_load_4d_2sm(dst + hb * (kHalfBytes / 2), tm, &bar->kv_full[stage], hb * kColsPerBox,
                                              g.seq.k_off + nb * kBlockN + rank * (kBlockN / 2), g.hk, g.seq.bk);
          } else {
            tma_load_4d_2sm(dst, tm, &bar->kv_full[stage], rank * kColsPerBox, g.seq.k_off + nb * kBlockN, g.hk, g.seq.bk);
          }
        }
      } else if (p.page_size == 0) {
        if (elect_one()) {
          mbar_arrive_expect_tx(&bar->kv_full[stage], boxes * kHalfBytes);
#pragma unroll
          for (int hb = 0; hb < kBoxesD; ++hb)
            if (hb < boxes) tma_load_4d(dst + hb * kHalfBytes, tm, &bar->kv_full[stage], hb * kColsPerBox,
                        g.seq.k_off + nb * kBlockN, g.hk, g.seq.bk);
        }
      } else {
        // paged: kv_box_rows = gcd(128, page_size) rows per box (a box never crosses a page); page ids come from the table
        const int32_t* tbl = p.page_table + int64_t(g.it.bi) * p.page_table_batch_stride;
        if (elect_one()) mbar_arrive_expect_tx(&bar->kv_full[stage], boxes * kHalfBytes);
        for (int r0 = 0; r0 < kBlockN; r0 += p.kv_box_rows) {
          const int row = nb * kBlockN + r0;
          const int page = tbl[min(row, max(g.seq.sk - 1, 0)) / p.page_size];  // rows past the end are masked; keep the table read in range
          if (elect_one()) {
#pragma unroll
            for (int hb = 0; hb < kBoxesD; ++hb)
              if (hb < boxes) tma_load_4d(dst + hb * kHalfBytes + r0 * kRowBytes, tm, &bar->kv_full[stage], hb * kColsPerBox,
                          row % p.page_size, g.hk, page);
          }
        }
      }
      __syncwarp();
      ring_advance<kRing>(ring);
    };
    if (!DUAL && kRing == 1 && !with_qv) {
      // One K and one V stage (16-bit hdim-256 instance): K(nb - 1) goes out BEFORE V(nb).  This warp waits in program order, and
      // V(nb)'s stage is free only when P V(nb + 1) has completed -- a full key block after QK(nb) released the K stage; behind
      // V(nb) the K load started that late and every key block paid load latency + both products in series.  The QK and PV warps
      // watch their own stages only, so the order between the kinds is free (not with qv: there QK(nb) also reads V(nb)).
      load_tile(0, g.n_hi - 1);
      for (int nb = g.n_hi - 1; nb >= g.n_lo; --nb) {
        if (nb - 1 >= g.n_lo) load_tile(0, nb - 1);
        load_tile(1, nb);
      }
    } else if (!DUAL) {
      // ring order == the order the MMA warp consumes: K(first), then V(nb), K(nb-1) per key block
      for (int nb = g.n_hi - 1; nb >= g.n_lo; --nb) { load_tile(0, nb); load_tile(1, nb); }
    } else {
      // two interleaved streams: K of both tiles' first blocks, then per step V_0, K_0(next), V_1, K_1(next)
      load_tile(0, g.n_hi - 1);
      if (g.n_hi - 2 >= g.n_lo) load_tile(0, g.n_hi - 2);
      for (int nb = g.n_hi - 1; nb >= g.n_lo; nb -= 2) {
        load_tile(1, nb);
        if (nb - 2 >= g.n_lo) load_tile(0, nb - 2);
        if (nb - 1 >= g.n_lo) load_tile(1, nb - 1);
        if (nb - 3 >= g.n_lo) load_tile(0, nb - 3);
      }
    }
  }
}

// ----------------------------------------------------------------------------- MMA issuers
// TMEM columns: S @0 (one score buffer shared by both Q tiles), P_0 @128, P_1 @192 (operand dtype), O_0 @256, O_1 @384.
//
// The two Q tiles take turns in the single S buffer.  A softmax warpgroup copies its scores into registers right after
// s_full[t] and arrives on s_free; from then on S belongs to the other tile, so S_t(next) = Q_t K^T never waits for
// P_t V.  P_t has its own columns; pv_done[t] tells the softmax warps that P_t (and O_t, for the rare rescale) may be
// touched again.
//
// Two warps feed the tensor pipe: the QK warp issues S = Q_t K^T (SS form) and owns the K stages of the ring, the PV
// warp issues O_t += P_t V (TS form, P in TMEM) and owns the V stages.  With a single issuing warp every barrier it
// polls (~110 clk even when complete) and every hand-off it waits for leaves the pipe idle -- measured 500 clk of
// 1520 per half key-block period; with two, one warp's waiting overlaps the other's issuing.  Both warps walk the
// same ring sequence (kIsQK selects which stages a warp acts on): K(first), then V(nb), K(nb-1) per key block, or in
// DUAL mode K_0, K_1, then V_0, K_0(next), V_1, K_1(next) per step.  All 32 lanes run the control flow and the waits
// (descriptors stay in uniform registers); one elected lane issues.
template <int HD, int ELT, bool FP8, bool DUAL, bool kIsQK, bool PAIR = false, bool RING = false>
FA_DEVICE void mma_role(const FwdKernelParams& p, uint32_t fmt, uint8_t* sQ, uint8_t* sKV, FwdBarriers* bar, uint32_t tmem) {
  using L = FwdSmem<HD, ELT, PAIR>;
  constexpr int kHalfBytes = kBlockN * 128;
  constexpr int kColsPerBox = 128 / ELT;
  constexpr int kKPerMma = 32 / ELT;                          // 16 (16-bit) / 32 (8-bit)
  constexpr int kSlicesPerBox = kColsPerBox / kKPerMma;       // 4
  constexpr int kQKSlices = HD / kKPerMma;
  constexpr int kPVSlices = kBlockN / kKPerMma;
  constexpr int kM = PAIR ? 2 * kBlockM : kBlockM;  // PAIR: one instruction covers the t-th tiles of both CTAs
  const uint32_t idesc = kIsQK ? make_idesc(fmt, fmt, kM, kBlockN, 0, 0) : make_idesc(fmt, fmt, kM, p.pv_n, 0, 1);
  // K-major operand tiles [128 rows][128 B] SW128: 8-row groups 1024 B apart (SBO); LBO unused (1).
  // V as MN-major B operand: 128-byte rows along dv, 8 key rows per swizzle atom (SBO = 1024 B), the next
  // 64 (16-bit) / 128 (8-bit) dv columns in the next box (LBO = one [128][128 B] region).
  constexpr uint32_t kHi = sdesc_hi(1024);
  constexpr uint32_t kLboK = sdesc_lo_lbo(16), kLboV = sdesc_lo_lbo(kHalfBytes);
  const uint32_t q_lo = (smem_u32(sQ) >> 4) | kLboK;
  const uint32_t kv_lo = smem_u32(sKV) >> 4;

  // completion signals: PAIR multicasts every arrival to the same barrier in both CTAs (the peer's producer and softmax
  // warps wait on their own copies)
  auto commit = [&](uint64_t* b) { if (PAIR) tc_commit_2sm(b); else tc_commit(b); };
  RingState ring;
  RingState ring_v_qk;  // qv: the QK warp also reads the V stages (score term Qv V^T); it only waits for them, the PV warp releases
  const bool with_qv = kIsQK && HD > 128 && p.qv_slices > 0;
  uint32_t q_it[kQTiles] = {0, 0}, p_it[kQTiles] = {0, 0}, o_it[kQTiles] = {0, 0};
  uint32_t qk_n = 0;  // QK batches issued so far == phases of s_free consumed + 1
  FA_PROF_DECL();     // 0 kv_full waits, 1 s_free waits, 2 p_full / o_empty waits, 3 issue, 4 everything else

  // S = Q_t K(stage)^T once the previous owner of S has taken its scores
  auto qk_op = [&](int t, int stage, bool last_use_of_q) {
    if (qk_n > 0) mbar_wait(&bar->s_free, (qk_n - 1) & 1, 0x48 + t);
    ++qk_n;
    int v_stage = 0;
    if (HD > 128 && with_qv) {
      // V(nb) must have landed before its K-major reading below; QK(nb) completes before softmax(nb), hence before the
      // PV(nb) whose commit releases the stage, so no release is needed here
      v_stage = L::kStages / 2 + ring_v_qk.stage;
      mbar_wait(&bar->kv_full[v_stage], ring_v_qk.phase, 0x43);
      ring_advance<L::kStages / 2>(ring_v_qk);
    }
    tc_fence_after();
    FA_PROF(1);
    if (elect_one()) {
      // qv: the Q tile sits in box 1 of the (single) K stage, the Q slot holds Qv
      const uint32_t a_lo = (HD > 128 && with_qv) ? (((smem_u32(sKV) + kHalfBytes) >> 4) | kLboK) : q_lo + t * (L::kQBytes >> 4);
      const uint32_t b_lo = (kv_lo + stage * (L::kKVBytes >> 4)) | kLboK;
#pragma unroll
      for (int ks = 0; ks < kQKSlices; ++ks) {
        const uint32_t off = (ks / kSlicesPerBox) * (kHalfBytes >> 4) + (ks % kSlicesPerBox) * 2;
        // PAIR: this CTA's K boxes hold 64 keys, so the next 64 head-dim columns are half as far away
        const uint32_t off_b = PAIR ? (ks / kSlicesPerBox) * (kHalfBytes >> 5) + (ks % kSlicesPerBox) * 2 : off;
        if (ks < p.qk_slices) {
          if (PAIR) mma_ss_2sm(tmem, a_lo + off, b_lo + off_b, kHi, idesc, ks > 0);
          else mma_ss<FP8>(tmem, a_lo + off, b_lo + off, kHi, idesc, ks > 0);
        }
      }
      if (HD > 128 && with_qv) {
        // S += Qv V^T: A = the Qv tile (Q slot), B = the V tile read K-major (rows = keys, 64 dv columns per box: the layout of
        // a K tile of head dim dv), contraction over dv
        const uint32_t v_lo = (kv_lo + v_stage * (L::kKVBytes >> 4)) | kLboK;
#pragma unroll
        for (int ks = 0; ks < kQKSlices; ++ks) {
          const uint32_t off = (ks / kSlicesPerBox) * (kHalfBytes >> 4) + (ks % kSlicesPerBox) * 2;
          if (ks < p.qv_slices) mma_ss<FP8>(tmem, q_lo + off, v_lo + off, kHi, idesc, 1u);
        }
      }
      commit(&bar->s_full[t]);
      if (last_use_of_q) commit(&bar->q_empty[t]);
      FA_LAT_STAMP(&bar->ts[t]);
    }
    __syncwarp();
    FA_PROF(3);
  };
  // O_t (+)= P_t V(stage)
  auto pv_op = [&](int t, int stage, bool first, bool last, int valid_rows) {
    // Key rows of the block past the end of the sequence are real memory (the tail of a dense cache, the rest of a page),
    // not TMA zero fill: their probabilities are 0, but 0 x NaN / Inf from an uninitialised cache would poison every row
    // of the tile.  Zero them in the V stage before the product (once per work item at most; PAIR: the peer's half is out
    // of reach of this warp, the variant requires finite cache tails).
    if (!PAIR && valid_rows < kBlockN) {
      uint8_t* vs = sKV + stage * L::kKVBytes;
      const int r0 = max(valid_rows, 0);
      for (int i = r0 * 8 + int(threadIdx.x & 31); i < kBlockN * 8; i += 32)
        for (int hb = 0; hb < p.v_boxes; ++hb) *reinterpret_cast<uint4*>(vs + hb * kHalfBytes + i * 16) = make_uint4(0u, 0u, 0u, 0u);
      fence_proxy_async_smem();
      __syncwarp();
    }
    if (first) mbar_wait(&bar->o_empty[t], (o_it[t] & 1) ^ 1, 0x50 + t);  // epilogue of the previous work item
    mbar_wait(&bar->p_full[t], p_it[t] & 1, 0x60 + t);
    ++p_it[t];
    tc_fence_after();
    FA_PROF(2);
    FA_LAT_ACC(6, &bar->ts[4 + t]);
    if (elect_one()) {
      const uint32_t b_lo = (kv_lo + stage * (L::kKVBytes >> 4)) | kLboV;
#pragma unroll
      for (int ks = 0; ks < kPVSlices; ++ks) {
          if (PAIR) mma_ts_2sm(tmem + 256 + t * 128, tmem + 128 + t * 64 + ks * 8, b_lo + ks * (kKPerMma * 128 >> 4), kHi, idesc, (!first || ks > 0));
          else mma_ts<FP8>(tmem + 256 + t * 128, tmem + 128 + t * 64 + ks * 8, b_lo + ks * (kKPerMma * 128 >> 4), kHi, idesc, (!first || ks > 0));
        }
      commit(&bar->pv_done[t]);
      if (last) commit(&bar->o_full[t]);
    }
    __syncwarp();
    FA_PROF(3);
    if (last) ++o_it[t];
  };
  // the next tile of this warp's kind (K tiles live in stages [0, kRing), V tiles in [kRing, 2 kRing)); -1 for the other kind
  constexpr int kRing = L::kStages / 2;
  auto next_stage = [&](bool is_k) {
    if (is_k != kIsQK) return -1;
    const int stage = (kIsQK ? 0 : kRing) + ring.stage;
    FA_PROF(4);
    mbar_wait(&bar->kv_full[stage], ring.phase, is_k ? 0x42 : 0x41);
    FA_PROF(0);
    ring_advance<kRing>(ring);
    return stage;
  };
  auto release = [&](int stage) {
    if (elect_one()) commit(&bar->kv_empty[stage]);
    __syncwarp();
  };

  const int n_work = total_work(p);
  for (int wi = 0;; ++wi) {
    const int w = RING ? 0 : work_index<PAIR>(wi);
    if (!RING && w >= n_work) break;
    int n_min[kQTiles], n_max[kQTiles], n_lo, n_hi, seq_k;
    bool end_of_list;
    {
      const WorkGeom g = item_geom<DUAL, (HD > 128), PAIR, RING>(p, bar, w, wi, 0);  // the MMA warps run in the pair's rank 0
      end_of_list = RING && g.w < 0;
      n_lo = g.n_lo; n_hi = g.n_hi; seq_k = g.seq.sk;
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) { n_min[t] = g.tr[t].n_min; n_max[t] = g.tr[t].n_max; }
    }
    item_geom_release<RING>(bar, wi);
    if (end_of_list) break;
    if (n_hi <= n_lo) continue;
    if (kIsQK) {
#pragma unroll
      for (int t = 0; t < kQTiles; ++t)
        if (n_max[t] > n_min[t]) { mbar_wait(&bar->q_full[t], q_it[t] & 1, 0x30 + t); ++q_it[t]; }
    }
    if constexpr (DUAL) {
      // tile t walks blocks n_max[t]-1, n_max[t]-3, ...
      int left[kQTiles];  // key blocks tile t still has to score
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) left[t] = (n_max[t] - n_min[t] + 1) / 2;
#pragma unroll
      for (int t = 0; t < kQTiles; ++t)
        if (left[t] > 0) {
          const int sk = next_stage(true);
          if (kIsQK) { qk_op(t, sk, left[t] == 1); release(sk); }
          --left[t];
        }
      const int steps = (n_hi - n_lo + 1) / 2;  // tile 0 never has fewer blocks than tile 1
      for (int j = 0; j < steps; ++j) {
#pragma unroll
        for (int t = 0; t < kQTiles; ++t) {
          const int nb = n_max[t] - 1 - 2 * j;
          if (nb < n_min[t]) continue;
          const int sv = next_stage(false);
          if (left[t] > 0) {
            const int sk = next_stage(true);
            if (kIsQK) { qk_op(t, sk, left[t] == 1); release(sk); }
            --left[t];
          }
          if (!kIsQK) { pv_op(t, sv, j == 0, nb - 2 < n_min[t], seq_k - nb * kBlockN); release(sv); }
        }
      }
    } else {
      {  // scores of the first (right-most) key block of the work item
        const int sk = next_stage(true);
        const int nb = n_hi - 1;
        if (kIsQK) {
#pragma unroll
          for (int t = 0; t < kQTiles; ++t)
            if (nb >= n_min[t] && nb < n_max[t]) qk_op(t, sk, n_min[t] == nb);
          release(sk);
        }
      }
      for (int nb = n_hi - 1; nb >= n_lo; --nb) {
        const bool has_next = nb - 1 >= n_lo;
        const int sv = next_stage(false);
        const int sk = has_next ? next_stage(true) : -1;
        if (kIsQK) {
          if (has_next) {
#pragma unroll
            for (int t = 0; t < kQTiles; ++t)
              if (nb - 1 >= n_min[t] && nb - 1 < n_max[t]) qk_op(t, sk, n_min[t] == nb - 1);
            release(sk);
          }
        } else {
#pragma unroll
          for (int t = 0; t < kQTiles; ++t)
            if (nb >= n_min[t] && nb < n_max[t]) pv_op(t, sv, nb == n_max[t] - 1, nb == n_min[t], seq_k - nb * kBlockN);
          release(sv);
        }
      }
    }
  }
  FA_PROF(4);
  FA_PROF_FLUSH();
}

}  // namespace fa100
