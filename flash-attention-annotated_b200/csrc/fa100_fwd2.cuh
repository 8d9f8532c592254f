// fa100_fwd2.cuh -- the prefill mainloop (head dims <= 128, two Q tiles per CTA): 16 warps in four roles.
//
// Same job as fa100_fwd_softmax.cuh's kernel (reference: hopper/mainloop_fwd_sm90_tma_gmma_ws.hpp:955-1352, softmax.h:92-168,
// epilogue_fwd.hpp:215-481), re-cut so that neither the tensor chain nor the softmax chain waits on the other tile:
//
//   warps 0-3 / 4-7   softmax of Q tile 0 / 1: one thread = one query row; S -> registers, row max, exponentials, P -> TMEM.
//                     Never touches O: a needed rescale is only announced (one float per row through shared memory).
//   warps 8-11        correction warpgroup: rescales O_t in tensor memory when a row max jumped, and owns the whole epilogue
//                     (normalise, convert, TMA / vector stores), so a softmax warp goes straight on to its next work item.
//   warp 12           the ONE MMA issuer.  Issue order per key block n: PV_0(n), QK_0(n-1), PV_1(n), QK_1(n-1).  tcgen05.mma
//                     instructions of one thread execute in order, so S_t(n-1) may overwrite the columns P_t(n) lives in
//                     without any hand-shake (P_t overlays the upper half of S_t), and the only things this warp ever waits
//                     for are P_t and the K / V stages.  While it waits for P_1 the tensor pipe drains PV_0 + QK_0.
//   warp 13           TMA producer (unchanged: producer_role).      warps 14-15 idle (complete the fourth warpgroup).
//
// TMEM: S_0 @0, S_1 @128 (fp32 scores), P_t @ S_t + 64 (operand dtype: 64 columns at 16 bit, 32 at e4m3), O_0 @256, O_1 @384.
// P is published in two pieces (3/4, then the rest) so that PV_t starts while the last exponentials are still in flight.
// Registers (setmaxnreg): softmax 184, correction 80, the rest 64  (8*32*184 + 4*32*80 + 4*32*64 = 65536).
#pragma once
#include "fa100_fwd_softmax.cuh"

namespace fa100 {

constexpr int kThreads2 = 16 * 32;
#ifndef FA100_REGS2
#define FA100_REGS2 184, 80, 64
#endif
constexpr int kRegs2[3] = {FA100_REGS2};
constexpr int kRegsSoftmax2 = kRegs2[0], kRegsCorr2 = kRegs2[1], kRegsOther2 = kRegs2[2];
static_assert(8 * kRegsSoftmax2 + 4 * kRegsCorr2 + 4 * kRegsOther2 <= 16 * 128, "register pool of 512 threads x 128");
constexpr int kCorrWarp0 = 8, kMmaWarp2 = 12, kLoadWarp2 = 13;

struct FwdBarriers2 {
  uint64_t q_full[kQTiles], q_empty[kQTiles];
  uint64_t s_full[kQTiles];    // MMA warp -> softmax t: S_t holds the scores of the next key block
  uint64_t p_full[kQTiles];    // softmax t (first 3/4 of P_t written) + correction (O_t rescaled / free) -> MMA warp
  uint64_t p_full2[kQTiles];   // softmax t: the last quarter of P_t is written
  uint64_t msg_full[kQTiles];  // softmax t -> correction: a message (rescale factor, or the final weights) is in sMsg
  uint64_t msg_empty[kQTiles]; // correction -> softmax t: the message has been read
  uint64_t o_full[kQTiles];    // MMA warp -> correction: the last PV of the work item has completed
  uint64_t kv_full[8], kv_empty[8];
  uint32_t tmem_base;
  uint32_t pad_[3];
};

// Shared memory of the v2 kernel: the v1 layout (Q tiles, K/V ring, O staging) + one message slot per row and tile.
template <int HD, int ELT>
struct FwdSmem2 {
  using L = FwdSmem<HD, ELT, false>;
  static constexpr int kMsgBytes = kQTiles * kBlockM * 2 * 4;  // two floats per row and tile
  static constexpr int kBarrierBytes = 512;                    // mbarriers + TMEM base (+ the trace build's counters in the last 128 B)
  static constexpr int kBudget = 227 * 1024 - kBarrierBytes - kMsgBytes;
  static constexpr int kStagesRaw = ((kBudget - L::kQSlots * L::kQBytes - L::kOBytes) / L::kKVBytes) & ~1;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static_assert(kStages >= 2 && kStages == L::kStages, "the producer role sizes its rings from FwdSmem");
  static constexpr int kDataBytes = L::kDataBytes;
  static constexpr int kTotalBytes = kDataBytes + kMsgBytes + kBarrierBytes;
  static_assert(sizeof(FwdBarriers2) <= kBarrierBytes - 128, "barrier block outgrew its slot");
  static_assert(kTotalBytes <= 227 * 1024, "shared memory over the 227 KB limit");
};

// ----------------------------------------------------------------------------- MMA issuer (one warp)
template <int HD, int ELT, bool FP8>
FA_DEVICE void mma2_role(const FwdKernelParams& p, uint32_t fmt, uint8_t* sQ, uint8_t* sKV, FwdBarriers2* bar, uint32_t tmem) {
  using L = FwdSmem<HD, ELT, false>;
  constexpr int kHalfBytes = kBlockN * 128;
  constexpr int kColsPerBox = 128 / ELT;
  constexpr int kKPerMma = 32 / ELT;                     // 16 (16-bit) / 32 (8-bit)
  constexpr int kSlicesPerBox = kColsPerBox / kKPerMma;  // 4
  constexpr int kQKSlices = HD / kKPerMma;
  constexpr int kPVSlices = kBlockN / kKPerMma;          // 8 / 4
  constexpr int kPVFirst = kPVSlices * 3 / 4;            // k-slices covered by the first piece of P
  constexpr int kRing = L::kStages / 2;
  const uint32_t idesc_qk = make_idesc(fmt, fmt, kBlockM, kBlockN, 0, 0);
  const uint32_t idesc_pv = make_idesc(fmt, fmt, kBlockM, p.pv_n, 0, 1);
  constexpr uint32_t kHi = sdesc_hi(1024);
  constexpr uint32_t kLboK = sdesc_lo_lbo(16), kLboV = sdesc_lo_lbo(kHalfBytes);
  const uint32_t q_lo = (smem_u32(sQ) >> 4) | kLboK;
  const uint32_t kv_lo = smem_u32(sKV) >> 4;
  RingState ring_k, ring_v;
  uint32_t q_it[kQTiles] = {0, 0}, p_it[kQTiles] = {0, 0};
  FA_PROF_DECL();  // 0 kv_full waits, 2 p_full waits, 3 issue, 4 everything else

  auto wait_k = [&]() {
    const int stage = ring_k.stage;
    FA_PROF(4);
    mbar_wait(&bar->kv_full[stage], ring_k.phase, 0x42);
    FA_PROF(0);
    ring_advance<kRing>(ring_k);
    return stage;
  };
  auto wait_v = [&]() {
    const int stage = kRing + ring_v.stage;
    FA_PROF(4);
    mbar_wait(&bar->kv_full[stage], ring_v.phase, 0x41);
    FA_PROF(0);
    ring_advance<kRing>(ring_v);
    return stage;
  };
  auto release = [&](int stage) {
    if (elect_one()) tc_commit(&bar->kv_empty[stage]);
    __syncwarp();
  };
  // S_t = Q_t K(stage)^T.  No wait: every earlier MMA that read P_t (the columns S_t shares) was issued before this one.
  auto qk_op = [&](int t, int stage, bool last_use_of_q) {
    tc_fence_after();
    if (elect_one()) {
      const uint32_t a_lo = q_lo + t * (L::kQBytes >> 4);
      const uint32_t b_lo = (kv_lo + stage * (L::kKVBytes >> 4)) | kLboK;
#pragma unroll
      for (int ks = 0; ks < kQKSlices; ++ks) {
        const uint32_t off = (ks / kSlicesPerBox) * (kHalfBytes >> 4) + (ks % kSlicesPerBox) * 2;
        if (ks < p.qk_slices) mma_ss<FP8>(tmem + t * 128, a_lo + off, b_lo + off, kHi, idesc_qk, ks > 0);
      }
      tc_commit(&bar->s_full[t]);
      if (last_use_of_q) tc_commit(&bar->q_empty[t]);
    }
    __syncwarp();
    FA_PROF(3);
  };
  // O_t (+)= P_t V(stage), in two pieces as P_t is published
  auto pv_op = [&](int t, int stage, bool first, bool last, int valid_rows) {
    if (valid_rows < kBlockN) {  // key rows past the end of the sequence are real memory: zero them (see fa100_fwd_roles.cuh pv_op)
      uint8_t* vs = sKV + stage * L::kKVBytes;
      const int r0 = max(valid_rows, 0);
      for (int i = r0 * 8 + int(threadIdx.x & 31); i < kBlockN * 8; i += 32)
        for (int hb = 0; hb < p.v_boxes; ++hb) *reinterpret_cast<uint4*>(vs + hb * kHalfBytes + i * 16) = make_uint4(0u, 0u, 0u, 0u);
      fence_proxy_async_smem();
      __syncwarp();
    }
    const uint32_t b_lo = (kv_lo + stage * (L::kKVBytes >> 4)) | kLboV;
    const uint32_t tP = tmem + t * 128 + 64, tO = tmem + 256 + t * 128;
    FA_PROF(4);
    mbar_wait(&bar->p_full[t], p_it[t] & 1, 0x60 + t);
    tc_fence_after();
    FA_PROF(2);
    if (elect_one()) {
#pragma unroll
      for (int ks = 0; ks < kPVFirst; ++ks)
        mma_ts<FP8>(tO, tP + ks * 8, b_lo + ks * (kKPerMma * 128 >> 4), kHi, idesc_pv, (!first || ks > 0));
    }
    __syncwarp();
    FA_PROF(3);
    mbar_wait(&bar->p_full2[t], p_it[t] & 1, 0x68 + t);
    tc_fence_after();
    FA_PROF(2);
    ++p_it[t];
    if (elect_one()) {
#pragma unroll
      for (int ks = kPVFirst; ks < kPVSlices; ++ks)
        mma_ts<FP8>(tO, tP + ks * 8, b_lo + ks * (kKPerMma * 128 >> 4), kHi, idesc_pv, 1u);
      if (last) tc_commit(&bar->o_full[t]);
    }
    __syncwarp();
    FA_PROF(3);
  };

  const int n_work = total_work(p);
  for (int wi = 0, w; (w = work_index<false>(wi)) < n_work; ++wi) {
    int n_min[kQTiles], n_max[kQTiles], n_lo, n_hi, seq_k;
    {
      const WorkGeom g = make_geom<false, false, false>(p, w, 0);
      n_lo = g.n_lo; n_hi = g.n_hi; seq_k = g.seq.sk;
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) { n_min[t] = g.tr[t].n_min; n_max[t] = g.tr[t].n_max; }
    }
    if (n_hi <= n_lo) continue;
#pragma unroll
    for (int t = 0; t < kQTiles; ++t)
      if (n_max[t] > n_min[t]) { mbar_wait(&bar->q_full[t], q_it[t] & 1, 0x30 + t); ++q_it[t]; }
    {  // scores of the first (right-most) key block
      const int sk = wait_k();
      const int nb = n_hi - 1;
#pragma unroll
      for (int t = 0; t < kQTiles; ++t)
        if (nb >= n_min[t] && nb < n_max[t]) qk_op(t, sk, n_min[t] == nb);
      release(sk);
    }
    for (int nb = n_hi - 1; nb >= n_lo; --nb) {
      const bool has_next = nb - 1 >= n_lo;
      const int sv = wait_v();
      int sk = -1;
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) {
        if (nb >= n_min[t] && nb < n_max[t]) pv_op(t, sv, nb == n_max[t] - 1, nb == n_min[t], seq_k - nb * kBlockN);
        if (has_next && nb - 1 >= n_min[t] && nb - 1 < n_max[t]) {
          if (sk < 0) sk = wait_k();
          qk_op(t, sk, n_min[t] == nb - 1);
        }
      }
      release(sv);
      if (has_next) {
        if (sk < 0) sk = wait_k();  // the producer loads K(nb - 1) whether or not a tile uses it
        release(sk);
      }
    }
  }
  FA_PROF(4);
  FA_PROF_FLUSH();
}

// ----------------------------------------------------------------------------- softmax (two warpgroups, one per Q tile)
// Per key block: wait S_t, copy it into registers, mask (boundary blocks only), row max, decide on a (lazy) rescale and
// announce it to the correction warpgroup, then the order-pinned exp pipeline (fa100_fwd_softmax.cuh) with the P stores
// folded in: chunk c of P goes to tensor memory as soon as its 16 pairs are packed, p_full[t] is signalled once 3/4 of the
// row is visible and p_full2[t] after the rest.  At the end of a work item the thread computes its row's final weights and
// LSE (hopper/softmax.h:137-155), stores the LSE, and hands the weights to the correction warpgroup.
#ifndef FA100_PUBLISH_AT
#define FA100_PUBLISH_AT 52  // pair index (0..63, in pack order) after which the first 3/4 of P is published
#endif

template <int HD, int ELT, bool FP8, bool F16>
FA_DEVICE void softmax2_role(const FwdKernelParams& p, float* sMsg, FwdBarriers2* bar, uint32_t tmem, int warp, int lane) {
  const int t = warp >> 2, q = warp & 3;
  const int row = q * 32 + lane;
  const uint32_t tS = tmem + (uint32_t(q * 32) << 16) + t * 128;
  const uint32_t tP = tS + 64;
  float* msg = sMsg + (t * kBlockM + row) * 2;
  uint32_t s_it = 0, msg_n = 0;  // msg_n: messages sent so far
  FA_PROF_DECL();  // 0 s_full wait, 1 S load, 2 mask + row max + message, 4 exp pipeline up to the first publish, 5 rest of it, 6 publish, 7 rest

  const int n_work = total_work(p);
  for (int wi = 0, w; (w = work_index<false>(wi)) < n_work; ++wi) {
    const RowGeom rg = row_geom<FP8, false, false, false>(p, make_geom<false, false, false>(p, w, 0), t, row, 0);
    if (!rg.has_rows || rg.n_max <= rg.n_min) continue;  // a tile without keys is written by the correction warpgroup alone
    const float scale_log2 = rg.scale_log2;
    const bool softcap = rg.softcap_pre != 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    for (int nb = rg.n_max - 1; nb >= rg.n_min; --nb) {
      FA_PROF(7);
      // this block will announce a rescale factor: its slot is free once the previous message has been read (long ago --
      // the correction warpgroup reads before it lets the previous PV_t go); settle that before the scores arrive.  The
      // first block of a work item sends nothing, so a softmax warp never waits for the previous item's epilogue here.
      if (msg_n > 0 && nb != rg.n_max - 1) mbar_wait(&bar->msg_empty[t], (msg_n - 1) & 1, 0x90 + t);
      mbar_wait(&bar->s_full[t], s_it & 1, 0x70 + t);
      FA_PROF(0);
      ++s_it;
      tc_fence_after();
      uint32_t s[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, s[c]);
      tmem_ld_wait();
      FA_PROF(1);
      const int n0 = nb * kBlockN;
      // boundary block for any row of this warp (warp-uniform): softcap first, then the mask (hopper/mainloop...:1073-1077)
      const bool boundary = __any_sync(0xffffffffu, softcap || n0 + kBlockN > rg.col_hi || n0 < rg.col_lo);
      if (boundary) {
        if (softcap) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(fast_tanh(__uint_as_float(s[c][i]) * rg.softcap_pre));
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int col = n0 + c * 32 + i;
            if (col >= rg.col_hi || col < rg.col_lo) s[c][i] = 0xff800000u;  // -inf
          }
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(s[c][i])); mx1 = fmaxf(mx1, __uint_as_float(s[c][i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(s[c][i + 2])); mx3 = fmaxf(mx3, __uint_as_float(s[c][i + 3]));
        }
      const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      const float m_cand = fmaxf(m_run, m_blk);
      // lazy rescale: keep the stale max unless it grew by more than 2^8 (NaN from -inf - -inf compares false)
      const bool grow = (m_cand - m_run) * scale_log2 > (FP8 ? kFp8RescaleThreshold : kRescaleThreshold);
      const float acc_scale = grow ? exp2f((m_run - m_cand) * scale_log2) : 1.0f;
      if (grow) m_run = m_cand;
      if (nb != rg.n_max - 1) {  // the first block has nothing to rescale; the correction warpgroup knows
        msg[0] = acc_scale;
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->msg_full[t]);
        ++msg_n;
      }
      const float m_sub = (m_run == -INFINITY) ? 0.f : m_run * scale_log2 - (FP8 ? kFp8Offset : 0.0f);
      FA_PROF(2);
      float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
      {
        constexpr int kAhead = 2, kBehind = 3;
        uint32_t pk[FP8 ? 32 : 64];
        uint16_t h8[FP8 ? 64 : 1];
        const float neg_m = -m_sub;
#pragma unroll
        for (int k = -kAhead; k < 64 + kBehind; ++k) {
          if (k + kAhead < 64) { const int a = k + kAhead; ffma2_v(s[a >> 4][(a & 15) * 2], s[a >> 4][(a & 15) * 2 + 1], scale_log2, neg_m); }
          if (k >= 0 && k < 64) {
            if (FA100_E2E_MOD > 0 && k < 48 && (k % (FA100_E2E_MOD > 0 ? FA100_E2E_MOD : 1)) == (FA100_E2E_MOD > 0 ? FA100_E2E_MOD : 1) - 1)
              ex2_poly2_v(s[k >> 4][(k & 15) * 2], s[k >> 4][(k & 15) * 2 + 1]);
            else { ex2_v(s[k >> 4][(k & 15) * 2]); ex2_v(s[k >> 4][(k & 15) * 2 + 1]); }
          }
          if (k - kBehind >= 0) {
            const int c = k - kBehind;
            const uint32_t p0 = s[c >> 4][(c & 15) * 2], p1 = s[c >> 4][(c & 15) * 2 + 1];
            if (c & 1) fadd2_v(sum2, sum3, p0, p1); else fadd2_v(sum0, sum1, p0, p1);
            if (!FP8) pk[c] = pack16_v<F16>(p0, p1);
            else {
              h8[c] = pack8_v(p0, p1);
              if (c & 1) pk[c >> 1] = uint32_t(h8[c - 1]) | (uint32_t(h8[c]) << 16);
            }
            // keys [32 j, 32 j + 32) of the block are complete: columns [16 j, +16) (16-bit) or [8 j, +8) (e4m3) of P_t
            if ((c & 15) == 15) {
              const int j = c >> 4;
              if (!FP8) tmem_st16(tP + j * 16, pk + j * 16); else tmem_st8(tP + j * 8, pk + j * 8);
            }
            if (c == FA100_PUBLISH_AT) {
              FA_PROF(4);
              tmem_st_wait();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&bar->p_full[t]);
              FA_PROF(6);
            }
          }
        }
      }
      FA_PROF(5);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->p_full2[t]);
      FA_PROF(6);
      l_run = l_run * acc_scale + ((sum0 + sum1) + (sum2 + sum3));
    }

    // ---- end of the work item: LSE and the weights of the final normalisation (hopper/softmax.h:137-155)
    {
      const WorkGeom g = make_geom<false, false, false>(p, w, 0);
      const bool split = p.o_accum != nullptr;
      const int part = g.it.split;
      const int in_tile = p.pack > 1 ? row / p.pack : row;
      const int pos = tile_pos0<false, false>(p, g.it, t, 0) + in_tile;
      const int hq = p.pack > 1 ? g.it.head * p.qheads_per_kv + row % p.pack : g.it.head;
      const bool row_valid = in_tile < p.pos_per_tile && pos < g.seq.sq;
      float out_scale = 1.0f;
      if (FP8) out_scale = p.v_descale ? p.v_descale[g.it.bi * p.v_descale_batch_stride + g.hk * p.v_descale_head_stride] : 1.f;
      float l_tot = l_run;
      const bool dead_row = (l_tot == 0.f) || (l_tot != l_tot);
      const float inv = dead_row ? 0.f : out_scale / l_tot;  // O carries the same 2^kFp8Offset as the row sum (fp8)
      if (FP8) l_tot *= exp2f(-kFp8Offset);
      const float lse_val = dead_row ? -INFINITY : (m_run * scale_log2 + log2f(l_tot)) * kLn2;
      const int64_t row_idx = g.seq.q_off + pos;
      float w_prev = 0.f, w_cur = inv;
      const bool merging = split && p.merge != 0;
      if (merging && row_valid) {  // fused merge with the running partial (hopper/flash_fwd_combine_kernel.h:345-458)
        const int64_t li = g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx;
        const float lse_prev = p.lse_accum[li];
        const float mx = fmaxf(lse_prev, lse_val);
        float lse_new = -INFINITY;
        if (mx != -INFINITY) {
          const float e_prev = __expf(lse_prev - mx), e_cur = __expf(lse_val - mx);
          lse_new = mx + __logf(e_prev + e_cur);
          w_prev = __expf(lse_prev - lse_new);
          w_cur = __expf(lse_val - lse_new) * inv;
        } else w_cur = 0.f;
        if (p.merge == 2) p.lse[li] = lse_new; else p.lse_accum[li] = lse_new;
      } else if (row_valid) {
        if (!split) p.lse[g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = lse_val;
        else p.lse_accum[part * p.lseacc_split_stride + g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = lse_val;
      }
      if (msg_n > 0) mbar_wait(&bar->msg_empty[t], (msg_n - 1) & 1, 0x90 + t);
      msg[0] = w_prev; msg[1] = w_cur;
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->msg_full[t]);
      ++msg_n;
    }
  }
  FA_PROF(7);
  FA_PROF_FLUSH();
}

// ----------------------------------------------------------------------------- correction warpgroup (warps 8-11)
// Walks the key blocks in the MMA warp's order.  For the first block of a tile it only says "O_t is free" (its own previous
// epilogue is done); for every later block it reads the row's rescale factor, multiplies O_t in tensor memory when any row
// of the warp needs it (rare: lazy rescale), and arrives on p_full[t].  No wait on the tensor pipe is needed: the factor of
// block n was computed from S_t(n), which the pipe produced after PV_t(n + 1).  At the end it takes the final weights
// (w_prev, w_cur) and runs the epilogue of hopper/epilogue_fwd.hpp:215-481: O leaves through a swizzled shared-memory
// box and TMA stores (dense batches), or as 16-byte vectors / fp32 partials / merged partials.
FA_DEVICE void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;\n" ::: "memory"); }

template <int HD, int ELT, bool FP8, bool F16>
FA_DEVICE void correction2_role(const FwdKernelParams& p, const CUtensorMap* tmO, uint8_t* sO, const float* sMsg, FwdBarriers2* bar,
                                uint32_t tmem, int warp, int lane) {
  const int q = warp & 3;
  const int row = q * 32 + lane;
  const uint32_t tBase = tmem + (uint32_t(q * 32) << 16);
  uint32_t msg_it[kQTiles] = {0, 0}, o_it[kQTiles] = {0, 0};
  uint32_t stage_n = 0;  // staging boxes handed to the TMA store engine so far (two boxes alternate)
  const bool issuer = (q == 0 && lane == 0);
  FA_PROF_DECL();  // 0 message waits, 1 rescale, 2 o_full wait, 3 epilogue, 7 rest

  const int n_work = total_work(p);
  for (int wi = 0, w; (w = work_index<false>(wi)) < n_work; ++wi) {
    const WorkGeom g = make_geom<false, false, false>(p, w, 0);
    for (int nb = g.n_hi - 1; nb >= g.n_lo; --nb) {
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) {
        if (!tile_uses(g, t, nb)) continue;
        if (nb != g.tr[t].n_max - 1) {
          FA_PROF(7);
          mbar_wait(&bar->msg_full[t], msg_it[t] & 1, 0xA0 + t);
          ++msg_it[t];
          const float scale = sMsg[(t * kBlockM + row) * 2];
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar->msg_empty[t]);
          FA_PROF(0);
          if (__any_sync(0xffffffffu, scale != 1.0f)) {
            tc_fence_after();
            const uint32_t tO = tBase + 256 + t * 128;
#pragma unroll
            for (int c = 0; c < HD / 32; ++c) {
              if (c * 32 >= p.dv) break;
              uint32_t o[32];
              tmem_ld32(tO + c * 32, o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * scale);
              tmem_st32(tO + c * 32, o);
            }
            tmem_st_wait();
            FA_PROF(1);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->p_full[t]);
      }
    }
    // ---- epilogue of both tiles
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      if (!g.tr[t].has_rows) continue;
      const bool live = tile_live(g, t);
      const uint32_t tO = tBase + 256 + t * 128;
      float w_prev = 0.f, w_cur = 0.f;
      FA_PROF(7);
      if (live) {
        mbar_wait(&bar->msg_full[t], msg_it[t] & 1, 0xA0 + t);
        ++msg_it[t];
        w_prev = sMsg[(t * kBlockM + row) * 2];
        w_cur = sMsg[(t * kBlockM + row) * 2 + 1];
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->msg_empty[t]);
        FA_PROF(0);
        mbar_wait(&bar->o_full[t], o_it[t] & 1, 0x80 + t);
        ++o_it[t];
        tc_fence_after();
        FA_PROF(2);
      }
      const bool split = p.o_accum != nullptr;
      const int part = g.it.split;
      const int in_tile = p.pack > 1 ? row / p.pack : row;
      const int pos = tile_pos0<false, false>(p, g.it, t, 0) + in_tile;
      const int hq = p.pack > 1 ? g.it.head * p.qheads_per_kv + row % p.pack : g.it.head;
      const bool row_valid = in_tile < p.pos_per_tile && pos < g.seq.sq;
      const int64_t row_idx = g.seq.q_off + pos;
      const bool merging = split && p.merge != 0;
      const bool to_out = merging && p.merge == 2;
      if (!live && row_valid) {
        // no key at all for this tile (e.g. causal with seqlen_k < seqlen_q): zeros and LSE = -inf; a merge keeps the stored partial
        if (merging) { w_prev = 1.f; if (to_out) p.lse[g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = p.lse_accum[g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx]; }
        else if (!split) p.lse[g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = -INFINITY;
        else p.lse_accum[part * p.lseacc_split_stride + g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = -INFINITY;
      }
      uint4* o16 = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.o) + g.seq.bq * p.o_batch_stride +
                                            row_idx * p.o_row_stride + hq * p.o_head_stride);
      float4* o32 = reinterpret_cast<float4*>(p.o_accum + part * p.oacc_split_stride + g.it.bi * p.oacc_batch_stride +
                                              hq * p.oacc_head_stride + row_idx * p.dv);
      if (p.o_tma) {
        // per 64-column slice: convert the row, write it into a [128][128 B] staging box (128-byte swizzle), one thread sends
        // the box with a bulk tensor store; rows past seqlen_q are clipped by the tensor map.  Two boxes alternate, so a box
        // is reused once the store before the previous one has read its data (wait_group.read 1 by the issuing thread).
#pragma unroll
        for (int sl = 0; sl < HD / 64; ++sl) {
          if (sl * 64 >= p.dv) break;
          uint8_t* stage = sO + (stage_n & 1) * FwdSmem<HD, ELT, false>::kOStageBytes;
          ++stage_n;
          if (issuer) tma_store_wait_read1();
          named_bar_sync(1, 128);
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            uint32_t o[32];
            if (live) { tmem_ld32(tO + sl * 64 + hf * 32, o); tmem_ld_wait(); }
            else {
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = 0u;
            }
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              uint4 u;
              uint32_t* up = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float a = __uint_as_float(o[ch * 8 + 2 * i]) * w_cur, b2 = __uint_as_float(o[ch * 8 + 2 * i + 1]) * w_cur;
                up[i] = F16 ? pack_f16(a, b2) : pack_bf16(a, b2);
              }
              *reinterpret_cast<uint4*>(stage + row * 128 + (((hf * 4 + ch) ^ (row & 7)) << 4)) = u;
            }
          }
          fence_proxy_async_smem();
          named_bar_sync(1, 128);
          if (issuer) {
            tma_store_4d(tmO, stage, sl * 64, pos - in_tile, hq, g.seq.bq);
            tma_store_commit();
          }
        }
      } else {
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
          if (c * 32 >= p.dv) break;
          uint32_t o[32];
          if (live) { tmem_ld32(tO + c * 32, o); tmem_ld_wait(); }
          else {
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = 0u;
          }
          if (row_valid && merging) {
#pragma unroll
            for (int v4 = 0; v4 < 8; ++v4) {
              if (c * 32 + v4 * 4 < p.dv) {
                const float4 pr = o32[c * 8 + v4];
                const float4 mg = make_float4(w_prev * pr.x + w_cur * __uint_as_float(o[v4 * 4]), w_prev * pr.y + w_cur * __uint_as_float(o[v4 * 4 + 1]),
                                              w_prev * pr.z + w_cur * __uint_as_float(o[v4 * 4 + 2]), w_prev * pr.w + w_cur * __uint_as_float(o[v4 * 4 + 3]));
                if (!to_out) o32[c * 8 + v4] = mg;
                else {
                  uint2 u;
                  u.x = F16 ? pack_f16(mg.x, mg.y) : pack_bf16(mg.x, mg.y);
                  u.y = F16 ? pack_f16(mg.z, mg.w) : pack_bf16(mg.z, mg.w);
                  reinterpret_cast<uint2*>(o16)[c * 8 + v4] = u;
                }
              }
            }
          } else if (row_valid) {
            if (!split) {
#pragma unroll
              for (int v8 = 0; v8 < 4; ++v8) {
                if (c * 32 + v8 * 8 < p.dv) {
                  uint4 u;
                  uint32_t* up = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float a = __uint_as_float(o[v8 * 8 + 2 * i]) * w_cur, b2 = __uint_as_float(o[v8 * 8 + 2 * i + 1]) * w_cur;
                    up[i] = F16 ? pack_f16(a, b2) : pack_bf16(a, b2);
                  }
                  o16[c * 4 + v8] = u;
                }
              }
            } else {
#pragma unroll
              for (int v4 = 0; v4 < 8; ++v4) {
                if (c * 32 + v4 * 4 < p.dv)
                  o32[c * 8 + v4] = make_float4(__uint_as_float(o[v4 * 4]) * w_cur, __uint_as_float(o[v4 * 4 + 1]) * w_cur,
                                                __uint_as_float(o[v4 * 4 + 2]) * w_cur, __uint_as_float(o[v4 * 4 + 3]) * w_cur);
              }
            }
          }
        }
      }
      FA_PROF(3);
      // O_t has been read: the "O_t is free" arrival for the next work item's first block follows in program order
      tc_fence_before();
    }
  }
  if (p.o_tma && issuer) tma_store_wait0();  // the staging boxes must outlive the last store
  FA_PROF(7);
  FA_PROF_FLUSH();
}

// ----------------------------------------------------------------------------- kernel entry
template <int HD, int ELT, bool FP8, bool F16>
__global__ void __launch_bounds__(kThreads2, 1)
fa100_fwd2_kernel(const __grid_constant__ FwdKernelParams p, const __grid_constant__ CUtensorMap tmQ,
                  const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                  const __grid_constant__ CUtensorMap tmO, uint32_t fmt) {
  using L = FwdSmem<HD, ELT, false>;
  using L2 = FwdSmem2<HD, ELT>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + L::kQSlots * L::kQBytes;
  uint8_t* sO = sKV + L::kStages * L::kKVBytes;
  float* sMsg = reinterpret_cast<float*>(smem + L2::kDataBytes);
  FwdBarriers2* bar = reinterpret_cast<FwdBarriers2*>(smem + L2::kDataBytes + L2::kMsgBytes);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  FA_TRACE_BEGIN();
  if (smem_u32(smem) & 1023u) __trap();
  if (warp == kLoadWarp2 && lane == 0) { tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); if (p.o_tma) tma_prefetch_desc(&tmO); }
  if (warp == kMmaWarp2 && lane == 0) {
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      mbar_init(&bar->q_full[t], 1); mbar_init(&bar->q_empty[t], 1);
      mbar_init(&bar->s_full[t], 1); mbar_init(&bar->p_full[t], 8); mbar_init(&bar->p_full2[t], 4);
      mbar_init(&bar->msg_full[t], 4); mbar_init(&bar->msg_empty[t], 4); mbar_init(&bar->o_full[t], 1);
    }
    for (int s = 0; s < 8; ++s) { mbar_init(&bar->kv_full[s], 1); mbar_init(&bar->kv_empty[s], 1); }
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc(&bar->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  pdl_wait();  // PDL: global memory is read only from here on (see fa100_ptx.cuh)
  pdl_launch_dependents();
  if (p.sched && !sched_matches(p)) {
    if (threadIdx.x == 0 && blockIdx.x == 0) { g_fa100_hang_info[0] = 0xDEAD0000u | 0x5CEDu; g_fa100_hang_info[1] = 0; g_fa100_hang_info[2] = 0; g_fa100_hang_info[3] = 0; __threadfence(); }
    __trap();
  }


  // register pool at launch: 512 threads x 128; the correction warps give 48 each, the last warpgroup 80 each, to the softmax warps (+64 each)
  if (warp < kSoftmaxWarps) {
    setmaxnreg_inc<kRegsSoftmax2>();
    softmax2_role<HD, ELT, FP8, F16>(p, sMsg, bar, tmem, warp, lane);
  } else if (warp < kCorrWarp0 + 4) {
    setmaxnreg_dec<kRegsCorr2>();
    correction2_role<HD, ELT, FP8, F16>(p, &tmO, sO, sMsg, bar, tmem, warp, lane);
  } else {
    setmaxnreg_dec<kRegsOther2>();
    if (warp == kMmaWarp2) mma2_role<HD, ELT, FP8>(p, fmt, sQ, sKV, bar, tmem);
    if (warp == kLoadWarp2) producer_role<HD, ELT, false, false, FwdBarriers2>(p, &tmQ, &tmK, &tmV, sQ, sKV, bar, 0);
  }
  FA_TRACE_END();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace fa100
