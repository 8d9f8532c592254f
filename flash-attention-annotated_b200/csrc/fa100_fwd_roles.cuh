// fa100_fwd_roles.cuh -- producer (TMA) and MMA-issuer roles of the forward kernel.
#pragma once
#include "fa100_fwd_kernel.cuh"

namespace fa100 {

// Work-item geometry every role derives identically (no communication needed).
struct WorkGeom {
  WorkItem it;
  SeqInfo seq;
  TileRange tr[kQTiles];
  int n_lo, n_hi;  // union of the tiles' key-block ranges; n_hi <= n_lo means nothing to do
  int hk;          // kv head
};

FA_DEVICE WorkGeom make_geom(const FwdKernelParams& p, int w) {
  WorkGeom g;
  g.it = decode_work(p, w);
  g.seq = load_seq(p, g.it.bi);
  g.hk = p.pack > 1 ? g.it.head : g.it.head / p.qheads_per_kv;
  g.n_lo = 0x7fffffff; g.n_hi = 0;
#pragma unroll
  for (int t = 0; t < kQTiles; ++t) {
    g.tr[t] = tile_range(p, g.seq, (g.it.m_blk * kQTiles + t) * p.pos_per_tile, g.it.split);
    if (g.tr[t].n_max > g.tr[t].n_min) { g.n_lo = min(g.n_lo, g.tr[t].n_min); g.n_hi = max(g.n_hi, g.tr[t].n_max); }
  }
  return g;
}
FA_DEVICE bool tile_live(const WorkGeom& g, int t) { return g.tr[t].n_max > g.tr[t].n_min; }
FA_DEVICE bool tile_uses(const WorkGeom& g, int t, int nb) { return nb >= g.tr[t].n_min && nb < g.tr[t].n_max; }

struct RingState { int stage = 0; uint32_t phase = 0; };
template <int STAGES> FA_DEVICE void ring_advance(RingState& r) { if (++r.stage == STAGES) { r.stage = 0; r.phase ^= 1; } }

// ----------------------------------------------------------------------------- TMA producer
// The whole warp walks the schedule (warp-uniform control flow keeps descriptors and coordinates in uniform
// registers); one elected lane issues the bulk copies.
template <int HD, int ELT>
FA_DEVICE void producer_role(const FwdKernelParams& p, const CUtensorMap* tmQ, const CUtensorMap* tmK,
                             const CUtensorMap* tmV, uint8_t* sQ, uint8_t* sKV, FwdBarriers* bar) {
  using L = FwdSmem<HD, ELT>;
  constexpr int kRowBytes = 128;                       // 128B-swizzled rows
  constexpr int kColsPerBox = kRowBytes / ELT;         // 64 (16-bit) or 128 (8-bit)
  constexpr int kBoxesD = HD / kColsPerBox > 0 ? HD / kColsPerBox : 1;  // boxes along the head dim
  constexpr int kHalfBytes = kBlockN * kRowBytes;      // one [128 rows][128 B] region
  RingState ring;
  uint32_t q_it[kQTiles] = {0, 0};
  const uint32_t q_tx = uint32_t(HD * ELT) * p.pack * p.pos_per_tile;
  for (int w = blockIdx.x; w < p.num_work; w += gridDim.x) {
    const WorkGeom g = make_geom(p, w);
    if (g.n_hi <= g.n_lo) continue;
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      if (!tile_live(g, t)) continue;
      mbar_wait(&bar->q_empty[t], (q_it[t] & 1) ^ 1, 0x10 + t);
      const int pos0 = (g.it.m_blk * kQTiles + t) * p.pos_per_tile;
      if (elect_one()) {
        mbar_arrive_expect_tx(&bar->q_full[t], q_tx);
#pragma unroll
        for (int hb = 0; hb < kBoxesD; ++hb)
          tma_load_5d(sQ + t * L::kQBytes + hb * kHalfBytes, tmQ, &bar->q_full[t], hb * kColsPerBox, 0,
                      g.seq.q_off + pos0, g.it.head, g.seq.bq);
      }
      __syncwarp();
      ++q_it[t];
    }
    for (int nb = g.n_hi - 1; nb >= g.n_lo; --nb) {
#pragma unroll
      for (int kv = 0; kv < 2; ++kv) {  // K tile then V tile of this key block
        const CUtensorMap* tm = kv == 0 ? tmK : tmV;
        mbar_wait(&bar->kv_empty[ring.stage], ring.phase ^ 1, 0x20 + kv);
        uint8_t* dst = sKV + ring.stage * L::kKVBytes;
        if (p.page_size == 0) {
          if (elect_one()) {
            mbar_arrive_expect_tx(&bar->kv_full[ring.stage], L::kKVBytes);
#pragma unroll
            for (int hb = 0; hb < kBoxesD; ++hb)
              tma_load_4d(dst + hb * kHalfBytes, tm, &bar->kv_full[ring.stage], hb * kColsPerBox,
                          g.seq.k_off + nb * kBlockN, g.hk, g.seq.bk);
          }
        } else {
          // paged: kv_box_rows = min(128, page_size) rows per box; page ids come from the table
          const int32_t* tbl = p.page_table + int64_t(g.it.bi) * p.page_table_batch_stride;
          if (elect_one()) mbar_arrive_expect_tx(&bar->kv_full[ring.stage], L::kKVBytes);
          for (int r0 = 0; r0 < kBlockN; r0 += p.kv_box_rows) {
            const int row = nb * kBlockN + r0;
            const int page = tbl[row / p.page_size];
            if (elect_one()) {
#pragma unroll
              for (int hb = 0; hb < kBoxesD; ++hb)
                tma_load_4d(dst + hb * kHalfBytes + r0 * kRowBytes, tm, &bar->kv_full[ring.stage], hb * kColsPerBox,
                            row % p.page_size, g.hk, page);
            }
          }
        }
        __syncwarp();
        ring_advance<L::kStages>(ring);
      }
    }
  }
}

// ----------------------------------------------------------------------------- MMA issuer
// One warp issues for both Q tiles, always tile 0 then tile 1 of a key block: S_t = Q_t K^T (SS form) and
// O_t += P_t V (TS form, P in TMEM).  The fixed order keeps the two tiles one MMA batch apart, so their softmax phases
// (which share the MUFU pipe) overlap as little as possible; with one issuing warp per tile both tiles stall on the
// same K/V stage, start together and contend (measured: 4530 vs 4270 clk per key-block pair).  All 32 lanes run the
// loop and the barrier waits (warp-uniform control flow keeps descriptors in uniform registers); one elected lane issues.
// TMEM columns: S_t = t*128, P_t = S_t (16-bit, 64 columns), O_t = 256 + t*128.
template <int HD, int ELT, bool FP8>
FA_DEVICE void mma_role(const FwdKernelParams& p, uint32_t fmt, uint8_t* sQ, uint8_t* sKV, FwdBarriers* bar, uint32_t tmem) {
  using L = FwdSmem<HD, ELT>;
  constexpr int kHalfBytes = kBlockN * 128;
  constexpr int kColsPerBox = 128 / ELT;
  constexpr int kKPerMma = 32 / ELT;                          // 16 (16-bit) / 32 (8-bit)
  constexpr int kSlicesPerBox = kColsPerBox / kKPerMma;       // 4
  constexpr int kQKSlices = HD / kKPerMma;
  constexpr int kPVSlices = kBlockN / kKPerMma;
  const uint32_t idesc_qk = make_idesc(fmt, fmt, kBlockM, kBlockN, 0, 0);
  const uint32_t idesc_pv = make_idesc(fmt, fmt, kBlockM, HD, 0, 1);
  // K-major operand tiles [128 rows][128 B] SW128: 8-row groups 1024 B apart (SBO); LBO unused (1).
  // V as MN-major B operand: 128-byte rows along dv, 8 key rows per swizzle atom (SBO = 1024 B), the next
  // 64 (16-bit) / 128 (8-bit) dv columns in the next box (LBO = one [128][128 B] region).
  constexpr uint32_t kHi = sdesc_hi(1024);
  constexpr uint32_t kLboK = sdesc_lo_lbo(16), kLboV = sdesc_lo_lbo(kHalfBytes);
  const uint32_t q_lo = (smem_u32(sQ) >> 4) | kLboK;
  const uint32_t kv_lo = smem_u32(sKV) >> 4;

  RingState ring;
  uint32_t q_it[kQTiles] = {0, 0}, p_it[kQTiles] = {0, 0}, o_it[kQTiles] = {0, 0};
  FA_PROF_DECL();  // 0 kv_full waits, 1 p_full / o_empty waits, 2 issue, 3 everything else

  auto issue_qk = [&](int t, int stage) {  // elected lane only
    const uint32_t a_lo = q_lo + t * (L::kQBytes >> 4);
    const uint32_t b_lo = (kv_lo + stage * (L::kKVBytes >> 4)) | kLboK;
#pragma unroll
    for (int ks = 0; ks < kQKSlices; ++ks) {
      const uint32_t off = (ks / kSlicesPerBox) * (kHalfBytes >> 4) + (ks % kSlicesPerBox) * 2;
      mma_ss<FP8>(tmem + t * 128, a_lo + off, b_lo + off, kHi, idesc_qk, ks > 0);
    }
    tc_commit(&bar->s_full[t]);
  };

  for (int w = blockIdx.x; w < p.num_work; w += gridDim.x) {
    int n_min[kQTiles], n_max[kQTiles], n_lo, n_hi;
    {
      const WorkGeom g = make_geom(p, w);
      n_lo = g.n_lo; n_hi = g.n_hi;
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) { n_min[t] = g.tr[t].n_min; n_max[t] = g.tr[t].n_max; }
    }
    if (n_hi <= n_lo) continue;
#pragma unroll
    for (int t = 0; t < kQTiles; ++t)
      if (n_max[t] > n_min[t]) mbar_wait(&bar->q_full[t], q_it[t] & 1, 0x30 + t);
    // prologue: S_t = Q_t K^T for the first (right-most) key block of the work item
    {
      mbar_wait(&bar->kv_full[ring.stage], ring.phase, 0x40);
      tc_fence_after();
      const int nb = n_hi - 1;
      if (elect_one()) {
#pragma unroll
        for (int t = 0; t < kQTiles; ++t)
          if (nb >= n_min[t] && nb < n_max[t]) {
            issue_qk(t, ring.stage);
            if (n_min[t] == nb) tc_commit(&bar->q_empty[t]);
          }
        tc_commit(&bar->kv_empty[ring.stage]);
      }
      __syncwarp();
#pragma unroll
      for (int t = 0; t < kQTiles; ++t)
        if (nb >= n_min[t] && nb < n_max[t] && n_min[t] == nb) ++q_it[t];
      ring_advance<L::kStages>(ring);
    }
    for (int nb = n_hi - 1; nb >= n_lo; --nb) {
      const RingState rv = ring;  // V(nb)
      ring_advance<L::kStages>(ring);
      const RingState rk = ring;  // K(nb-1)
      const bool has_next = nb - 1 >= n_lo;
      FA_PROF(3);
      mbar_wait(&bar->kv_full[rv.stage], rv.phase, 0x41);
      if (has_next) { mbar_wait(&bar->kv_full[rk.stage], rk.phase, 0x42); ring_advance<L::kStages>(ring); }
      FA_PROF(0);
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) {
        const bool use_pv = nb >= n_min[t] && nb < n_max[t];
        const bool use_qk = has_next && nb - 1 >= n_min[t] && nb - 1 < n_max[t];
        const bool first = nb == n_max[t] - 1;
        if (use_pv) {
          if (first) mbar_wait(&bar->o_empty[t], (o_it[t] & 1) ^ 1, 0x50 + t);  // epilogue of the previous work item
          mbar_wait(&bar->p_full[t], p_it[t] & 1, 0x60 + t);
          ++p_it[t];
          tc_fence_after();
        }
        FA_PROF(1);
        if (elect_one()) {
          if (use_pv) {
            const uint32_t b_lo = (kv_lo + rv.stage * (L::kKVBytes >> 4)) | kLboV;
#pragma unroll
            for (int ks = 0; ks < kPVSlices; ++ks)
              mma_ts<FP8>(tmem + 256 + t * 128, tmem + t * 128 + ks * 8, b_lo + ks * (kKPerMma * 128 >> 4), kHi, idesc_pv, (!first || ks > 0));
            if (nb == n_min[t]) tc_commit(&bar->o_full[t]);
          }
          if (use_qk) {
            issue_qk(t, rk.stage);
            if (n_min[t] == nb - 1) tc_commit(&bar->q_empty[t]);
          }
          if (t == kQTiles - 1) {
            tc_commit(&bar->kv_empty[rv.stage]);
            if (has_next) tc_commit(&bar->kv_empty[rk.stage]);
          }
        }
        __syncwarp();
        FA_PROF(2);
        if (use_pv && nb == n_min[t]) ++o_it[t];
        if (use_qk && n_min[t] == nb - 1) ++q_it[t];
      }
    }
  }
  FA_PROF(3);
  FA_PROF_FLUSH();
}

}  // namespace fa100
