// fa100_fwd_softmax.cuh -- softmax / correction / epilogue role (one warpgroup per Q tile) and the kernel entry.
#pragma once
#include "fa100_fwd_roles.cuh"

namespace fa100 {

FA_DEVICE uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
FA_DEVICE uint32_t pack_f16(float lo, float hi) {
  __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
FA_DEVICE float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
FA_DEVICE float fast_tanh(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// Packed f32x2 arithmetic (one issue slot for two scores of the row), order-pinned for the exp pipeline: the compiler
// front end keeps volatile asm statements in program order, so
// the PTX carries the software pipeline written in softmax_role (scale-subtract two pairs ahead of the exponentials,
// row sum and pack three pairs behind them) instead of one burst of FFMA2s followed by stretches with one MUFU per
// five issue slots (what the scheduler made of the per-chunk loops: a lone warp ran at 12.5 clk per score, not 8).
FA_DEVICE void ffma2_v(uint32_t& a, uint32_t& b, float sc, float m) {
  asm volatile("{\n\t.reg .b64 x, s2, m2;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 s2, {%2, %2};\n\tmov.b64 m2, {%3, %3};\n\t"
               "fma.rn.ftz.f32x2 x, x, s2, m2;\n\tmov.b64 {%0, %1}, x;\n\t}" : "+r"(a), "+r"(b) : "f"(sc), "f"(m));
}
FA_DEVICE void ex2_v(uint32_t& x) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(x)); }
FA_DEVICE void fadd2_v(float& acc0, float& acc1, uint32_t a, uint32_t b) {
  asm volatile("{\n\t.reg .b64 x, y;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 y, {%2, %3};\n\tadd.rn.ftz.f32x2 x, x, y;\n\tmov.b64 {%0, %1}, x;\n\t}"
               : "+f"(acc0), "+f"(acc1) : "r"(a), "r"(b));
}
template <bool F16> FA_DEVICE uint32_t pack16_v(uint32_t lo, uint32_t hi) {
  uint32_t r;
  if (F16) asm volatile("cvt.rn.f16x2.f32 %0, %2, %1;" : "=r"(r) : "r"(lo), "r"(hi));
  else asm volatile("cvt.rn.bf16x2.f32 %0, %2, %1;" : "=r"(r) : "r"(lo), "r"(hi));
  return r;
}
FA_DEVICE uint16_t pack8_v(uint32_t lo, uint32_t hi) {
  uint16_t r;
  asm volatile("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(r) : "r"(lo), "r"(hi));
  return r;
}
// Polynomial exp2 on a share of the scores: every kMod-th score pair of the first kLim pairs of a row takes ex2_poly2_v (FMA
// pipe) instead of the MUFU.  On by default for 16-bit head dim 64 only (every third pair of all 64): there the tensor work per
// score is half that of head dim 128 and the MUFU is the one wall (+5 % causal, +8 % non-causal at 8 k tokens,
// profiles/r02q_ab_poly_exp2_hdim64.log); at head dim 128 and for e4m3 every setting measured neutral or slower
// (profiles/r02_ab_poly_exp2_default_kernel.log), so those stay build-time experiment switches, default off.
#ifndef FA100_E2E_MOD
#define FA100_E2E_MOD 0
#endif
#ifndef FA100_E2E_MOD_FP8
#define FA100_E2E_MOD_FP8 FA100_E2E_MOD
#endif
#ifndef FA100_E2E_MOD_D64
#define FA100_E2E_MOD_D64 3
#endif
#ifndef FA100_E2E_LIM
#define FA100_E2E_LIM 48
#endif
#ifndef FA100_E2E_LIM_D64
#define FA100_E2E_LIM_D64 64
#endif
// exp2 of two scores on the FMA pipe instead of the MUFU (Cody-Waite split + degree-3 polynomial on [0, 1), the scheme of
// the reference's Blackwell kernel, flash_attn/cute/utils.py:601-681): x <= 0 here, clamped at -127; relative error ~1e-4,
// far below the bf16 / e4m3 rounding of P that follows.
FA_DEVICE void ex2_poly2_v(uint32_t& a, uint32_t& b) {
  asm volatile(
      "{\n\t.reg .f32 x0, x1, c;\n\t.reg .b64 x, r, rb, f, pl, k;\n\t.reg .b32 r0, r1, p0, p1;\n\t"
      "mov.b32 x0, %0;\n\tmov.b32 x1, %1;\n\t"
      "max.ftz.f32 x0, x0, 0fC2FE0000;\n\tmax.ftz.f32 x1, x1, 0fC2FE0000;\n\t"
      "mov.b64 x, {x0, x1};\n\t"
      "mov.f32 c, 0f4B400000;\n\tmov.b64 k, {c, c};\n\t"        // 2^23 + 2^22
      "add.rm.ftz.f32x2 r, x, k;\n\t"                               // floor(x) in the low mantissa bits
      "sub.rn.ftz.f32x2 rb, r, k;\n\t"
      "sub.rn.ftz.f32x2 f, x, rb;\n\t"                              // fractional part in [0, 1)
      "mov.f32 c, 0f3D9DF09D;\n\tmov.b64 pl, {c, c};\n\t"         // 0.077119089663
      "mov.f32 c, 0f3E6906A4;\n\tmov.b64 k, {c, c};\n\t"          // 0.227564394474
      "fma.rn.ftz.f32x2 pl, pl, f, k;\n\t"
      "mov.f32 c, 0f3F31F519;\n\tmov.b64 k, {c, c};\n\t"          // 0.695146143436
      "fma.rn.ftz.f32x2 pl, pl, f, k;\n\t"
      "mov.f32 c, 0f3F800000;\n\tmov.b64 k, {c, c};\n\t"          // 1.0
      "fma.rn.ftz.f32x2 pl, pl, f, k;\n\t"
      "mov.b64 {r0, r1}, r;\n\tmov.b64 {p0, p1}, pl;\n\t"
      "shl.b32 r0, r0, 23;\n\tshl.b32 r1, r1, 23;\n\t"
      "add.s32 %0, r0, p0;\n\tadd.s32 %1, r1, p1;\n\t}"
      : "+r"(a), "+r"(b));
}

// What one softmax thread needs about its row while it walks the key blocks of a work item.  Kept this small on
// purpose: with 227 KB of shared memory carved out the L1 is tiny, so a register spill in the step loop costs an L2
// round trip; everything the epilogue needs is recomputed there instead of being carried across the loop.
struct RowGeom {
  int n_min, n_max;    // key blocks [n_min, n_max) of this Q tile
  int col_lo, col_hi;  // visible key columns [col_lo, col_hi) of this row
  float scale_log2;
  float softcap_pre;   // softmax_scale / softcap (x q_descale x k_descale for e4m3); 0 = no softcap
  bool has_rows;
};

template <bool FP8, bool DUAL, bool SOLO, bool PAIR = false>
FA_DEVICE RowGeom row_geom(const FwdKernelParams& p, const WorkGeom& g, int t, int row, int rank = 0) {
  RowGeom r;
  const TileRange tr = t == 0 ? g.tr[0] : g.tr[1];  // no dynamic indexing: that would put the struct in local memory
  r.has_rows = tr.has_rows;
  r.n_min = tr.n_min; r.n_max = tr.n_max;
  const int pos = tile_pos0<DUAL || SOLO, PAIR>(p, g.it, t, rank) + (p.pack > 1 ? row / p.pack : row);
  const int shift = g.seq.sk - g.seq.sq;
  int col_hi = g.seq.sk, col_lo = 0;
  if (p.is_causal) col_hi = min(col_hi, pos + 1 + shift);
  if (p.is_local) {
    if (p.window_right >= 0) col_hi = min(col_hi, pos + 1 + shift + p.window_right);
    if (p.window_left >= 0) col_lo = max(col_lo, pos + shift - p.window_left);
    if (p.attention_chunk > 0) {
      const int x = pos + shift, lo_c = x - floor_mod(x, p.attention_chunk);
      col_lo = max(col_lo, lo_c);
      col_hi = min(col_hi, lo_c + p.attention_chunk);
    }
  }
  r.col_lo = col_lo; r.col_hi = col_hi;
  r.scale_log2 = p.scale_log2;
  r.softcap_pre = p.softcap_pre;
  if (FP8) {
    // q and k descales fold into the softmax scale (hopper/flash_fwd_kernel_sm90.h:409-415); with softcap they fold into
    // the factor in front of tanh instead (the scores leave tanh in true units; same place the reference puts them)
    const float qd = p.q_descale ? p.q_descale[g.it.bi * p.q_descale_batch_stride + g.hk * p.q_descale_head_stride] : 1.f;
    const float kd = p.k_descale ? p.k_descale[g.it.bi * p.k_descale_batch_stride + g.hk * p.k_descale_head_stride] : 1.f;
    if (p.softcap_pre != 0.f) r.softcap_pre *= qd * kd; else r.scale_log2 *= qd * kd;
  }
  return r;
}

// Online softmax follows hopper/softmax.h:92-168 (exp2 with the scale folded in, fp32 running max / sum,
// LSE = max*scale + ln(sum)) with Blackwell's row-per-lane TMEM layout and lazy rescaling: softmax warp w (0..7) works
// on Q tile t = w / 4 and owns TMEM lanes (= query rows) 32*(w%4)..; one thread owns one row, so the row max and the
// row sum need no shuffles.  The step loop has two bodies: a lean one for interior key blocks (no mask, no softcap; it
// is the instruction stream tools/ubench/ubench_softmax.cu measures in isolation) and a general one for boundary blocks.
// -DFA100_PROF_EPI (cycle-bucket build only): the key loop's marks all go to bucket 0 and buckets 1-6 split the time between
// key loops: 1 loop exit + geometry + LSE, 2 wait o_full, 3 O: TMEM -> registers -> convert, 4 staging barriers + stores,
// 5 o_empty arrive + loop back, 6 next item's row_geom; 7 stays "top of the first block".
#ifdef FA100_PROF_EPI
#define FA_PROF_L(i) FA_PROF(0)
#define FA_PROF_E(i) FA_PROF(i)
#undef FA_LAT_ACC
#define FA_LAT_ACC(i, word) do { } while (0)
#else
#define FA_PROF_L(i) FA_PROF(i)
#define FA_PROF_E(i) do { } while (0)
#endif
template <int HD, int ELT, bool FP8, bool F16, bool DUAL, bool PAIR = false, bool RING = false>
FA_DEVICE void softmax_role(const FwdKernelParams& p, const CUtensorMap* tmO, uint8_t* sO, FwdBarriers* bar, uint32_t tmem, int warp, int lane, int rank = 0) {
  const int t = warp >> 2, q = warp & 3;
  const int row = q * 32 + lane;
  const uint32_t tS = tmem + (uint32_t(q * 32) << 16);  // the score buffer both tiles take turns in
  const uint32_t tP = tS + 128 + t * 64;                // this tile's probabilities (operand dtype)
  const uint32_t tO = tS + 256 + t * 128;
  uint32_t s_it = 0, o_it = 0, pv_n = 0;                // pv_n: P blocks published so far == PV batches requested
  // the barriers the MMA warps wait on live in the pair's leader: the peer's softmax warps arrive there over the cluster
  const uint32_t a_s_free = PAIR ? mapa_u32(smem_u32(&bar->s_free), 0) : 0, a_p_full = PAIR ? mapa_u32(smem_u32(&bar->p_full[t]), 0) : 0,
                 a_o_empty = PAIR ? mapa_u32(smem_u32(&bar->o_empty[t]), 0) : 0;
  FA_PROF_DECL();  // 0 s_full wait, 1 S load, 2 s_free + mask + row max, 3 rescale, 4 exp chunks 0-1, 5 exp chunks 2-3, 6 st wait + publish, 7 rest

  const int n_work = total_work(p);
  for (int wi = 0;; ++wi) {
    int w = RING ? 0 : work_index<PAIR>(wi);
    if (!RING && w >= n_work) break;
    RowGeom rg;
    {
      const WorkGeom g0 = item_geom<DUAL, (HD > 128), PAIR, RING>(p, bar, w, wi, rank);
      if (RING && g0.w < 0) { item_geom_release<RING>(bar, wi); break; }
      w = g0.w;
      rg = row_geom<FP8, DUAL, (HD > 128), PAIR>(p, g0, t, row, rank);
    }
    FA_PROF_E(6);
    if (!rg.has_rows) { item_geom_release<RING>(bar, wi); continue; }
    const float scale_log2 = rg.scale_log2;
    const bool softcap = p.softcap_pre != 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    for (int nb = rg.n_max - 1; nb >= rg.n_min; nb -= (DUAL ? 2 : 1)) {
      FA_PROF_L(7);
      mbar_wait(&bar->s_full[t], s_it & 1, 0x70 + t);
      FA_PROF_L(0);
      FA_LAT_ACC(3, &bar->ts[t]);
      ++s_it;
      tc_fence_after();
      uint32_t s[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, s[c]);
      tmem_ld_wait();
      FA_PROF_L(1);
      // the scores are in registers: hand S to the other tile's QK
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (PAIR) mbar_arrive_cluster(a_s_free); else mbar_arrive(&bar->s_free); }
      if (q == 3) FA_LAT_STAMP(&bar->ts[2]);
      FA_TRACE_POINT(0xC4);
      // P_t and O_t may be touched again once the previous P_t V has drained; poll now, block only right before the
      // first write (normally long complete by then)
      bool pv_drained = pv_n == 0;
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
        // visible columns of this row inside the block: [lo, hi) with compile-time column indices on the other side of the
        // compare (one ISETP + one SEL per score; the left limit only when some row of the warp has one inside the block)
        const int hi = rg.col_hi - n0, lo = rg.col_lo - n0;
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i >= hi) s[c][i] = 0xff800000u;  // -inf
        if (__any_sync(0xffffffffu, lo > 0)) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c * 32 + i < lo) s[c][i] = 0xff800000u;
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
      FA_PROF_L(2);
      const float m_cand = fmaxf(m_run, m_blk);
      // lazy rescale: keep the stale max unless it grew by more than 2^8 (NaN from -inf - -inf compares false)
      const bool grow = (m_cand - m_run) * scale_log2 > (FP8 ? kFp8RescaleThreshold : kRescaleThreshold);
      const float acc_scale = grow ? exp2f((m_run - m_cand) * scale_log2) : 1.0f;
      if (grow) m_run = m_cand;
      if (nb != rg.n_max - 1 && __any_sync(0xffffffffu, grow)) {
        if (!pv_drained) { mbar_wait(&bar->pv_done[t], (pv_n - 1) & 1, 0x78 + t); pv_drained = true; }
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
          if (c * 32 >= p.dv) break;
          uint32_t o[32];
          tmem_ld32(tO + c * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * acc_scale);
          tmem_st32(tO + c * 32, o);
        }
      }
      const float m_sub = (m_run == -INFINITY) ? 0.f : m_run * scale_log2 - (FP8 ? kFp8Offset : 0.0f);
      FA_PROF_L(2);
      // P = exp2(S * scale - max), packed to the operand dtype and written over S: keys [32c, 32c + 32) of the tile go to
      // columns [16c, +16) (16-bit operands) or [8c, +8) (e4m3)
      float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
      {
        // One straight-line pipeline over the 64 score pairs of the row: pair k + 2 gets its scale-subtract while pair k
        // is in the MUFU pipe and pair k - 3 is summed and packed, so the exponentials issue at an even pace from the
        // first pair to the last.  P stays in the registers the scores leave behind and is written to tensor memory in
        // one go at the end (by then the previous P_t V has long drained, so the wait below is a formality).
        constexpr int kAhead = 2, kBehind = 3;
        uint32_t pk[FP8 ? 32 : 64];
        uint16_t h8[FP8 ? 64 : 1];
        const float neg_m = -m_sub;
#pragma unroll
        for (int k = -kAhead; k < 64 + kBehind; ++k) {
          if (k + kAhead < 64) { const int a = k + kAhead; ffma2_v(s[a >> 4][(a & 15) * 2], s[a >> 4][(a & 15) * 2 + 1], scale_log2, neg_m); }
          if (k >= 0 && k < 64) {
            constexpr bool kD64 = !FP8 && HD == 64;
            constexpr int kMod = FP8 ? FA100_E2E_MOD_FP8 : (kD64 ? FA100_E2E_MOD_D64 : FA100_E2E_MOD);
            constexpr int kLim = kD64 ? FA100_E2E_LIM_D64 : FA100_E2E_LIM;
            if (kMod > 0 && k < kLim && (k % (kMod > 0 ? kMod : 1)) == (kMod > 0 ? kMod : 1) - 1) ex2_poly2_v(s[k >> 4][(k & 15) * 2], s[k >> 4][(k & 15) * 2 + 1]);
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
          }
        }
        FA_PROF_L(4);
        if (!pv_drained) mbar_wait(&bar->pv_done[t], (pv_n - 1) & 1, 0x78 + t);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (!FP8) tmem_st16(tP + c * 16, pk + c * 16);
          else tmem_st8(tP + c * 8, pk + c * 8);
        }
      }
      FA_PROF_L(5);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (PAIR) mbar_arrive_cluster(a_p_full); else mbar_arrive(&bar->p_full[t]); }
      if (q == 3) FA_LAT_STAMP(&bar->ts[4 + t]);
      FA_TRACE_POINT(0xA0 + t);
      ++pv_n;
      FA_PROF_L(6);
      l_run = l_run * acc_scale + ((sum0 + sum1) + (sum2 + sum3));
    }

    // ---- epilogue: O / rowsum, LSE (hopper/softmax.h:137-155, epilogue_fwd.hpp:215-481); geometry is recomputed here
    {
      // RING: the slot is still this warp's (released below); otherwise recomputed rather than carried through the key loop
      const WorkGeom g = item_geom<DUAL, (HD > 128), PAIR, RING>(p, bar, w, wi, rank);
      const bool split = p.o_accum != nullptr;  // fp32 partials (split-KV, or one hop of ring attention)
      const int part = DUAL ? g.it.split * 2 + t : g.it.split;  // index of this partial
      const bool live = rg.n_max > rg.n_min;
      const int in_tile = p.pack > 1 ? row / p.pack : row;
      const int pos = tile_pos0<DUAL || (HD > 128), PAIR>(p, g.it, t, rank) + in_tile;
      const int hq = p.pack > 1 ? g.it.head * p.qheads_per_kv + row % p.pack : g.it.head;
      const bool row_valid = in_tile < p.pos_per_tile && pos < g.seq.sq;
      float out_scale = 1.0f;
      if (FP8) out_scale = p.v_descale ? p.v_descale[g.it.bi * p.v_descale_batch_stride + g.hk * p.v_descale_head_stride] : 1.f;
      float l_tot = l_run;
      const bool dead_row = (l_tot == 0.f) || (l_tot != l_tot);
      const float inv = dead_row ? 0.f : out_scale / l_tot;  // O carries the same 2^kFp8Offset as the row sum (fp8)
      if (FP8) l_tot *= exp2f(-kFp8Offset);                        // undo the offset for the LSE
      const float lse_val = dead_row ? -INFINITY : (m_run * scale_log2 + log2f(l_tot)) * kLn2;
      const int64_t row_idx = g.seq.q_off + pos;
      // Fused merge (ring attention): this call's (O_t, LSE_t) joins the running fp32 partial already in o_accum /
      // lse_accum with the combine arithmetic (hopper/flash_fwd_combine_kernel.h:345-458) instead of going through
      // memory and a separate combine launch.  w_prev scales the stored row, w_cur the row in tensor memory.
      float w_prev = 0.f, w_cur = inv;
      const bool merging = split && p.merge != 0;
      const bool to_out = merging && p.merge == 2;  // the last hop writes the output dtype
      if (merging && row_valid) {
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
        if (to_out) p.lse[li] = lse_new; else p.lse_accum[li] = lse_new;
      } else if (row_valid) {
        if (!split) p.lse[g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = lse_val;
        else p.lse_accum[part * p.lseacc_split_stride + g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = lse_val;
      }
      FA_PROF_E(1);
      if (live) { mbar_wait(&bar->o_full[t], o_it & 1, 0x80 + t); ++o_it; tc_fence_after(); }
      FA_PROF_E(2);
      uint4* o16 = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.o) + g.seq.bq * p.o_batch_stride +
                                            row_idx * p.o_row_stride + hq * p.o_head_stride);
      float4* o32 = reinterpret_cast<float4*>(p.o_accum + part * p.oacc_split_stride + g.it.bi * p.oacc_batch_stride +
                                              hq * p.oacc_head_stride + row_idx * p.dv);
      // packed (cu_seqlens_q) / seqused_q batches: a box that runs past the end of its sequence would write into the next one
      // (or into rows the caller keeps), so only tiles that lie wholly inside take the TMA path (tile-uniform test)
      if (HD <= 128 && p.o_tma && (!p.o_tma_check || pos - in_tile + kBlockM <= g.seq.sq)) {
        // TMA-store epilogue (role of the reference's smem_o + TMA store, hopper/epilogue_fwd.hpp:262-318): per 64-column
        // slice every thread converts its row, writes it into this tile's [128][128 B] staging box in the 128-byte
        // swizzle, and one thread sends the box with a bulk tensor store; rows past seqlen_q are clipped by the tensor map.
        // The staging box is free again once the previous store has read it (wait_group.read by the issuing thread,
        // at the top of the next slice or the next work item).
        uint8_t* stage = sO + t * FwdSmem<HD, ELT, PAIR>::kOStageBytes;
        const bool issuer = (q == 0 && lane == 0);
#pragma unroll
        for (int sl = 0; sl < HD / 64; ++sl) {
          if (sl * 64 >= p.dv) break;
          uint32_t pk[32];
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            uint32_t o[32];
            if (live) { tmem_ld32(tO + sl * 64 + hf * 32, o); tmem_ld_wait(); }
            else {
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = 0u;
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float a = __uint_as_float(o[2 * i]) * inv, b2 = __uint_as_float(o[2 * i + 1]) * inv;
              pk[hf * 16 + i] = F16 ? pack_f16(a, b2) : pack_bf16(a, b2);
            }
          }
          FA_PROF_E(3);
          if (issuer) tma_store_wait_read0();
          named_bar_sync(1 + t, 128);
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) {
            uint4 u = make_uint4(pk[ch * 4], pk[ch * 4 + 1], pk[ch * 4 + 2], pk[ch * 4 + 3]);
            *reinterpret_cast<uint4*>(stage + row * 128 + ((ch ^ (row & 7)) << 4)) = u;
          }
          fence_proxy_async_smem();
          named_bar_sync(1 + t, 128);
          if (issuer) {
            tma_store_4d(tmO, stage, sl * 64, g.seq.q_off + pos - in_tile, hq, g.seq.bq);
            tma_store_commit();
          }
          FA_PROF_E(4);
        }
      } else
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
                  const float a = __uint_as_float(o[v8 * 8 + 2 * i]) * inv, b2 = __uint_as_float(o[v8 * 8 + 2 * i + 1]) * inv;
                  up[i] = F16 ? pack_f16(a, b2) : pack_bf16(a, b2);
                }
                o16[c * 4 + v8] = u;
              }
            }
          } else {
#pragma unroll
            for (int v4 = 0; v4 < 8; ++v4) {
              if (c * 32 + v4 * 4 < p.dv)
                o32[c * 8 + v4] = make_float4(__uint_as_float(o[v4 * 4]) * inv, __uint_as_float(o[v4 * 4 + 1]) * inv,
                                              __uint_as_float(o[v4 * 4 + 2]) * inv, __uint_as_float(o[v4 * 4 + 3]) * inv);
            }
          }
        }
      }
      if (live) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (PAIR) mbar_arrive_cluster(a_o_empty); else mbar_arrive(&bar->o_empty[t]); }
      }
      FA_PROF_E(5);
    }
    item_geom_release<RING>(bar, wi);
  }
  if (HD <= 128 && p.o_tma && q == 0 && lane == 0) tma_store_wait0();  // the staging box must outlive the last store
  FA_PROF(7);
  FA_PROF_FLUSH();
}

// ----------------------------------------------------------------------------- kernel entry
template <int HD, int ELT, bool FP8, bool F16, bool DUAL, bool PAIR = false>
__global__ void __launch_bounds__(kThreads, 1)
fa100_fwd_kernel(const __grid_constant__ FwdKernelParams p, const __grid_constant__ CUtensorMap tmQ,
                 const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                 const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmQv, uint32_t fmt) {
  using L = FwdSmem<HD, ELT, PAIR>;
  extern __shared__ __align__(1024) uint8_t smem[];  // 128-byte swizzle atoms need 1 KB alignment
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + L::kQSlots * L::kQBytes;
  uint8_t* sO = sKV + L::kStages * L::kKVBytes;
  FwdBarriers* bar = reinterpret_cast<FwdBarriers*>(smem + L::kDataBytes);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int rank = PAIR ? int(cluster_ctarank()) : 0;  // PAIR: launched as clusters of two CTAs; rank 0 issues the MMAs
  FA_TRACE_BEGIN();
  if (smem_u32(smem) & 1023u) __trap();
  if (warp == kLoadWarp && lane == 0) { tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); if (p.o_tma) tma_prefetch_desc(&tmO); }
  if (warp == kMmaWarp && lane == 0) {
    // PAIR: the softmax warps of both CTAs arrive on the leader's s_free / p_full / o_empty
    constexpr int kSmxArrivals = PAIR ? 8 : 4;
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      mbar_init(&bar->q_full[t], 1); mbar_init(&bar->q_empty[t], 1);
      mbar_init(&bar->s_full[t], 1); mbar_init(&bar->p_full[t], kSmxArrivals); mbar_init(&bar->pv_done[t], 1);
      mbar_init(&bar->o_full[t], 1); mbar_init(&bar->o_empty[t], kSmxArrivals);
    }
    mbar_init(&bar->s_free, kSmxArrivals);
    for (int s = 0; s < 8; ++s) { mbar_init(&bar->kv_full[s], 1); mbar_init(&bar->kv_empty[s], 1); }
    // geometry ring: consumers are the eight softmax warps, the producer and (rank 0 of a pair only) the two issuing warps
    for (int s = 0; s < kGeomSlots; ++s) { mbar_init(&bar->geom_full[s], 1); mbar_init(&bar->geom_empty[s], (PAIR && rank != 0) ? 9 : 11); }
    fence_mbar_init();
  }
  if (warp == 0) {
    if (PAIR) { tmem_alloc_2sm(&bar->tmem_base, 512); tmem_relinquish_2sm(); }
    else { tmem_alloc(&bar->tmem_base, 512); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer's barriers exist before anything is sent to them
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  // PDL: everything above touched only shared / tensor memory and kernel parameters; global memory written by the kernels
  // in front of this one (schedule, append-KV, the caller's producers) is read only from here on
  pdl_wait();
  pdl_launch_dependents();
  if (p.sched && !sched_matches(p)) {
    // a schedule built for other sizes / num_splits / pack_gqa / sm_margin / kernel variant would index partials and
    // row blocks that do not exist: fail the launch loudly (tag 0x5C4ED in fa100_debug_hang_info)
    if (threadIdx.x == 0 && blockIdx.x == 0) { g_fa100_hang_info[0] = 0xDEAD0000u | 0x5CEDu; g_fa100_hang_info[1] = 0; g_fa100_hang_info[2] = 0; g_fa100_hang_info[3] = 0; __threadfence(); }
    __trap();
  }


  // register pool at launch: 384 threads x 168; warps 8-11 give 112 each to the 8 softmax warps (+56 each)
  if (warp < kSoftmaxWarps) {
    setmaxnreg_inc<224>();
    softmax_role<HD, ELT, FP8, F16, DUAL, PAIR, true>(p, &tmO, sO, bar, tmem, warp, lane, rank);
  } else {
    setmaxnreg_dec<56>();
    if (rank == 0) {
      if (warp == kMmaWarp) mma_role<HD, ELT, FP8, DUAL, true, PAIR, true>(p, fmt, sQ, sKV, bar, tmem);
      if (warp == kPvWarp) mma_role<HD, ELT, FP8, DUAL, false, PAIR, true>(p, fmt, sQ, sKV, bar, tmem);
    }
    if (warp == kLoadWarp) producer_role<HD, ELT, DUAL, PAIR, FwdBarriers, true>(p, &tmQ, &tmK, &tmV, sQ, sKV, bar, rank, &tmQv);
    if (warp == kGeomWarp) geom_role<DUAL, (HD > 128), PAIR>(p, bar, rank);
  }
  FA_TRACE_END();
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // nothing of the peer (barriers, K / V halves, tensor memory) may go away while the other CTA still uses it
  if (warp == 0) { if (PAIR) tmem_dealloc_2sm(tmem, 512); else tmem_dealloc(tmem, 512); }
}

}  // namespace fa100
