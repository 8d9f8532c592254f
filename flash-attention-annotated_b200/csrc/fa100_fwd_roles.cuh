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
// All 32 lanes run the loop and the barrier waits (warp-uniform), one elected lane issues tcgen05.mma / commit.
// TMEM columns: S_t = t*128, P_t = S_t (16-bit, 64 columns), O_t = 256 + t*128.
template <int HD, int ELT, bool FP8>
FA_DEVICE void mma_role(const FwdKernelParams& p, uint32_t fmt, uint8_t* sQ, uint8_t* sKV, FwdBarriers* bar,
                        uint32_t tmem) {
  using L = FwdSmem<HD, ELT>;
  constexpr int kHalfBytes = kBlockN * 128;
  constexpr int kColsPerBox = 128 / ELT;
  constexpr int kKPerMma = 32 / ELT;                          // 16 (16-bit) / 32 (8-bit)
  constexpr int kSlicesPerBox = kColsPerBox / kKPerMma;       // 4
  constexpr int kQKSlices = HD / kKPerMma;
  constexpr int kPVSlices = kBlockN / kKPerMma;
  const uint32_t idesc_qk = make_idesc(fmt, fmt, kBlockM, kBlockN, 0, 0);
  const uint32_t idesc_pv = make_idesc(fmt, fmt, kBlockM, HD, 0, 1);
  // K-major operand tiles [128 rows][128 B] SW128: 8-row groups 1024 B apart (SBO); LBO unused.
  constexpr uint64_t kDescK = make_sdesc_base(16, 1024);
  // V as MN-major B operand: 128-byte rows along dv, 8 key rows per swizzle atom (SBO = 1024 B),
  // next 64(16-bit)/128(8-bit) dv columns in the next box (LBO = one [128][128 B] region).
  constexpr uint64_t kDescV = make_sdesc_base(kHalfBytes, 1024);
  const uint32_t sQ_addr = smem_u32(sQ), sKV_addr = smem_u32(sKV);

  RingState ring;
  uint32_t q_it[kQTiles] = {0, 0}, p_it[kQTiles] = {0, 0}, o_it[kQTiles] = {0, 0};

  auto issue_qk = [&](int t, int stage) {
    const uint32_t a0 = sQ_addr + t * L::kQBytes, b0 = sKV_addr + stage * L::kKVBytes;
    if (elect_one()) {
#pragma unroll
      for (int ks = 0; ks < kQKSlices; ++ks) {
        const uint32_t off = (ks / kSlicesPerBox) * kHalfBytes + (ks % kSlicesPerBox) * 32;
        const uint64_t ad = sdesc_at(kDescK, a0 + off), bd = sdesc_at(kDescK, b0 + off);
        if (FP8) mma_f8_ss(tmem + t * 128, ad, bd, idesc_qk, ks > 0);
        else     mma_f16_ss(tmem + t * 128, ad, bd, idesc_qk, ks > 0);
      }
      tc_commit(&bar->s_full[t]);
    }
    __syncwarp();
  };
  auto commit = [&](uint64_t* b) { if (elect_one()) tc_commit(b); __syncwarp(); };
  auto issue_pv = [&](int t, int stage, bool accumulate) {
    const uint32_t b0 = sKV_addr + stage * L::kKVBytes;
    if (elect_one()) {
#pragma unroll
      for (int ks = 0; ks < kPVSlices; ++ks) {
        const uint64_t bd = sdesc_at(kDescV, b0 + ks * kKPerMma * 128);
        const uint32_t a_tmem = tmem + t * 128 + ks * 8;  // 8 columns (32 B of P) per K slice
        if (FP8) mma_f8_ts(tmem + 256 + t * 128, a_tmem, bd, idesc_pv, (accumulate || ks > 0));
        else     mma_f16_ts(tmem + 256 + t * 128, a_tmem, bd, idesc_pv, (accumulate || ks > 0));
      }
    }
    __syncwarp();
  };

  for (int w = blockIdx.x; w < p.num_work; w += gridDim.x) {
    const WorkGeom g = make_geom(p, w);
    if (g.n_hi <= g.n_lo) continue;
#pragma unroll
    for (int t = 0; t < kQTiles; ++t)
      if (tile_live(g, t)) { mbar_wait(&bar->q_full[t], q_it[t] & 1, 0x30 + t); }
    tc_fence_after();
    // prologue: S_t = Q_t K^T for the first (right-most) key block
    {
      mbar_wait(&bar->kv_full[ring.stage], ring.phase, 0x40);
      tc_fence_after();
#pragma unroll
      for (int t = 0; t < kQTiles; ++t)
        if (tile_uses(g, t, g.n_hi - 1)) {
          issue_qk(t, ring.stage);
          if (g.tr[t].n_min == g.n_hi - 1) { commit(&bar->q_empty[t]); ++q_it[t]; }
        }
      commit(&bar->kv_empty[ring.stage]);
      ring_advance<L::kStages>(ring);
    }
    for (int nb = g.n_hi - 1; nb >= g.n_lo; --nb) {
      const RingState rv = ring;  // V(nb)
      ring_advance<L::kStages>(ring);
      const RingState rk = ring;  // K(nb-1)
      const bool has_next = nb - 1 >= g.n_lo;
      mbar_wait(&bar->kv_full[rv.stage], rv.phase, 0x41);
      if (has_next) { mbar_wait(&bar->kv_full[rk.stage], rk.phase, 0x42); ring_advance<L::kStages>(ring); }
      tc_fence_after();
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) {
        if (tile_uses(g, t, nb)) {
          const bool first = nb == g.tr[t].n_max - 1;
          if (first) mbar_wait(&bar->o_empty[t], (o_it[t] & 1) ^ 1, 0x50 + t);  // epilogue of the previous item
          mbar_wait(&bar->p_full[t], p_it[t] & 1, 0x60 + t);
          ++p_it[t];
          tc_fence_after();
          FA_TRACE_POINT(0xB0 + t);
          issue_pv(t, rv.stage, !first);
          FA_TRACE_POINT(0xB2 + t);
          if (nb == g.tr[t].n_min) { commit(&bar->o_full[t]); ++o_it[t]; }
        }
        if (has_next && tile_uses(g, t, nb - 1)) {
          issue_qk(t, rk.stage);
          FA_TRACE_POINT(0xB4 + t);
          if (g.tr[t].n_min == nb - 1) { commit(&bar->q_empty[t]); ++q_it[t]; }
        }
      }
      commit(&bar->kv_empty[rv.stage]);
      if (has_next) commit(&bar->kv_empty[rk.stage]);
      FA_TRACE_POINT(0xB6);
    }
  }
}

}  // namespace fa100
