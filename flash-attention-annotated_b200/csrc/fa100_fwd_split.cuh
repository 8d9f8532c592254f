// fa100_fwd_split.cuh -- the prefill kernel with TWO softmax threads per query row (16 softmax warps, 640 threads).
//
// Same TMEM layout, barriers, MMA issuers and producer as fa100_fwd_kernel (two 128-row Q tiles, one score buffer they take
// turns in); what changes is who does the softmax arithmetic.  With one thread per row a scheduler hosts two softmax warps
// whose instruction streams are in order, and the MUFU (128 exponentials per row and key block, 8 clk per warp
// instruction) idles whenever both are in a non-exponential phase: 73 % busy at C2.  Here warp
// w and warp w + 8 share the TMEM lanes (= rows) of one quarter of a tile and each takes 64 of the 128 key columns, so a
// scheduler hosts four shorter streams.  Per key block the two threads of a row swap the maxima of their halves through
// shared memory (one fp32 slot each, a 64-thread named barrier), so both derive bit-identical running maxima and rescale
// decisions; each keeps the row sum of its own columns and rescales / stores its own half of the O columns, and the two
// partial sums meet once per work item the same way.  (First version: no swap, every thread read the other half's scores
// too and reduced them to a max -- S was handed to the other tile's QK ~350 clk late and the S ping-pong became the critical
// loop, profiles/r02g_c2_split_rows_phase_samples.txt.)
#pragma once
#include "fa100_fwd_softmax.cuh"

namespace fa100 {

constexpr int kSplitSoftmaxWarps = 16;
constexpr int kSplitThreads = 20 * 32;
constexpr int kSplitExchangeBytes = kQTiles * 2 * kBlockM * 4;  // one fp32 slot per (tile, half, row)

template <int HD, int ELT>
struct FwdSmemSplit {
  using L = FwdSmem<HD, ELT, false>;
  // two sets of slots used alternately let a swap get by with ONE barrier (a thread's next write goes to the other set; by
  // the time it returns to this one the peer has passed a later barrier, hence read it).  bf16 / fp16 at head dim 128 has
  // 2 KB left in the 227 KB: one set, and a second barrier after the read.
  static constexpr bool kDouble = L::kTotalBytes + 2 * kSplitExchangeBytes <= 227 * 1024;
  static constexpr int kXchBytes = (kDouble ? 2 : 1) * kSplitExchangeBytes;
  static constexpr int kTotalBytes = L::kTotalBytes + kXchBytes;
  static_assert(kTotalBytes <= 227 * 1024, "split kernel outgrew shared memory");
};

// raw (pre-softcap) maximum of 32 scores of the row at key columns [col0, col0 + 32), masked to [col_lo, col_hi) when asked
template <bool MASK>
FA_DEVICE float chunk_max(const uint32_t (&x)[32], int col0, int col_lo, int col_hi) {
  float a = -INFINITY, b = -INFINITY, c = -INFINITY, d = -INFINITY;
#pragma unroll
  for (int i = 0; i < 32; i += 4) {
    float v0 = __uint_as_float(x[i]), v1 = __uint_as_float(x[i + 1]), v2 = __uint_as_float(x[i + 2]), v3 = __uint_as_float(x[i + 3]);
    if (MASK) {
      const int col = col0 + i;
      if (col >= col_hi || col < col_lo) v0 = -INFINITY;
      if (col + 1 >= col_hi || col + 1 < col_lo) v1 = -INFINITY;
      if (col + 2 >= col_hi || col + 2 < col_lo) v2 = -INFINITY;
      if (col + 3 >= col_hi || col + 3 < col_lo) v3 = -INFINITY;
    }
    a = fmaxf(a, v0); b = fmaxf(b, v1); c = fmaxf(c, v2); d = fmaxf(d, v3);
  }
  return fmaxf(fmaxf(a, b), fmaxf(c, d));
}

// Softmax warp w of 16: half = w / 8 (key columns [64 half, +64) of every block), tile t = (w / 4) % 2, TMEM lane quarter w % 4.
// SOLO (head dims above 128, one Q tile per work item): the eight softmax warps the kernel has anyway -- four of them idle
// in the one-thread-per-row kernel -- share the single tile: half = w / 4, t = 0, the 224-register budget is untouched.
template <int HD, int ELT, bool FP8, bool F16, bool SOLO = false>
FA_DEVICE void softmax_split_role(const FwdKernelParams& p, const CUtensorMap* tmO, uint8_t* sO, float* xch, FwdBarriers* bar,
                                  uint32_t tmem, int warp, int lane) {
  const int half = SOLO ? warp >> 2 : warp >> 3, t = SOLO ? 0 : (warp >> 2) & 1, q = warp & 3;
  const int row = q * 32 + lane;
  const uint32_t tS = tmem + (uint32_t(q * 32) << 16);
  const uint32_t tSown = tS + half * 64;
  const uint32_t tP = tS + 128 + t * 64 + half * (FP8 ? 16 : 32);  // this thread's half of P_t (operand dtype)
  const uint32_t tO = tS + 256 + t * 128;
  constexpr int kOChunks = HD / 64;       // 32-column chunks of O per thread
  const int c0 = half * kOChunks;         // first of them
  constexpr bool kDouble = FwdSmemSplit<HD, ELT>::kDouble;
  const int my_slot = (t * 2 + half) * kBlockM + row, peer_slot = (t * 2 + (1 - half)) * kBlockM + row;
  const int pair_bar = 3 + t * 4 + q;     // named barrier of the two warps that share these rows (ids 1, 2: the tiles' epilogues)
  uint32_t s_it = 0, o_it = 0, pv_n = 0, xset = 0;
  // hand v to the row's other thread and get its value (see FwdSmemSplit::kDouble for the one- / two-barrier forms)
  auto pair_swap = [&](float v) -> float {
    float* slots = xch + (kDouble ? xset * (kSplitExchangeBytes / 4) : 0);
    slots[my_slot] = v;
    named_bar_sync(pair_bar, 64);
    const float o = slots[peer_slot];
    if (kDouble) xset ^= 1; else named_bar_sync(pair_bar, 64);
    return o;
  };
  FA_PROF_DECL();

  const int n_work = total_work(p);
  for (int wi = 0, w; (w = work_index<false>(wi)) < n_work; ++wi) {
    const RowGeom rg = row_geom<FP8, false, SOLO, false>(p, make_geom<false, SOLO, false>(p, w, 0), t, row);
    if (!rg.has_rows) continue;
    const float scale_log2 = rg.scale_log2;
    const bool softcap = p.softcap_pre != 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    for (int nb = rg.n_max - 1; nb >= rg.n_min; --nb) {
      FA_PROF(7);
      mbar_wait(&bar->s_full[t], s_it & 1, 0x70 + t);
      FA_PROF(0);
      ++s_it;
      tc_fence_after();
      const int n0 = nb * kBlockN;
      const bool boundary = __any_sync(0xffffffffu, softcap || n0 + kBlockN > rg.col_hi || n0 < rg.col_lo);
      uint32_t s[2][32];
      tmem_ld32(tSown, s[0]);
      tmem_ld32(tSown + 32, s[1]);
      tmem_ld_wait();
      FA_PROF(1);
      // every score this thread needs is in registers: hand S to the other tile's QK
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->s_free);
      bool pv_drained = pv_n == 0;
      const int own0 = n0 + half * 64;
      float m_raw;  // raw (pre-softcap), masked maximum: of this half, then of the row
      if (boundary) {
        m_raw = fmaxf(chunk_max<true>(s[0], own0, rg.col_lo, rg.col_hi), chunk_max<true>(s[1], own0 + 32, rg.col_lo, rg.col_hi));
        // softcap first, then the mask (hopper/mainloop...:1073-1077)
        if (softcap) {
#pragma unroll
          for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) s[c][i] = __float_as_uint(fast_tanh(__uint_as_float(s[c][i]) * rg.softcap_pre));
        }
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int col = own0 + c * 32 + i;
            if (col >= rg.col_hi || col < rg.col_lo) s[c][i] = 0xff800000u;  // -inf
          }
      } else {
        m_raw = fmaxf(chunk_max<false>(s[0], 0, 0, 0), chunk_max<false>(s[1], 0, 0, 0));
      }
      m_raw = fmaxf(m_raw, pair_swap(m_raw));
      // Both threads of the row now hold the same m_raw (max of the same two numbers) and apply the same function to it.
      // With softcap the reference point is tanh of the raw max: tanh.approx may differ
      // from the largest capped score by an ulp, which the lazy rescale (stale max, P up to 2^8) tolerates by design.
      const float m_blk = (softcap && m_raw != -INFINITY) ? fast_tanh(m_raw * rg.softcap_pre) : m_raw;
      FA_PROF(2);
      const float m_cand = fmaxf(m_run, m_blk);
      const bool grow = (m_cand - m_run) * scale_log2 > (FP8 ? kFp8RescaleThreshold : kRescaleThreshold);
      const float acc_scale = grow ? exp2f((m_run - m_cand) * scale_log2) : 1.0f;
      if (grow) m_run = m_cand;
      if (nb != rg.n_max - 1 && __any_sync(0xffffffffu, grow)) {
        if (!pv_drained) { mbar_wait(&bar->pv_done[t], (pv_n - 1) & 1, 0x78 + t); pv_drained = true; }
        tc_fence_after();
#pragma unroll
        for (int cc = 0; cc < kOChunks; ++cc) {
          const int c = c0 + cc;
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
      FA_PROF(3);
      float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
      {
        // the order-pinned exp pipeline of softmax_role over this thread's 32 score pairs
        constexpr int kPairs = 32, kAhead = 2, kBehind = 3;
        uint32_t pk[FP8 ? 16 : 32];
        uint16_t h8[FP8 ? 32 : 1];
        const float neg_m = -m_sub;
#pragma unroll
        for (int k = -kAhead; k < kPairs + kBehind; ++k) {
          if (k + kAhead < kPairs) { const int a = k + kAhead; ffma2_v(s[a >> 4][(a & 15) * 2], s[a >> 4][(a & 15) * 2 + 1], scale_log2, neg_m); }
          if (k >= 0 && k < kPairs) { ex2_v(s[k >> 4][(k & 15) * 2]); ex2_v(s[k >> 4][(k & 15) * 2 + 1]); }
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
        FA_PROF(4);
        if (!pv_drained) mbar_wait(&bar->pv_done[t], (pv_n - 1) & 1, 0x78 + t);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (!FP8) tmem_st16(tP + c * 16, pk + c * 16);
          else tmem_st8(tP + c * 8, pk + c * 8);
        }
      }
      FA_PROF(5);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->p_full[t]);
      ++pv_n;
      FA_PROF(6);
      l_run = l_run * acc_scale + ((sum0 + sum1) + (sum2 + sum3));
    }

    // ---- epilogue (softmax_role's, with each thread on its own half of the O columns)
    {
      const WorkGeom g = make_geom<false, SOLO, false>(p, w, 0);
      const bool split = p.o_accum != nullptr;
      const int part = g.it.split;
      const bool live = rg.n_max > rg.n_min;
      const int in_tile = p.pack > 1 ? row / p.pack : row;
      const int pos = tile_pos0<SOLO, false>(p, g.it, t, 0) + in_tile;
      const int hq = p.pack > 1 ? g.it.head * p.qheads_per_kv + row % p.pack : g.it.head;
      const bool row_valid = in_tile < p.pos_per_tile && pos < g.seq.sq;
      float out_scale = 1.0f;
      if (FP8) out_scale = p.v_descale ? p.v_descale[g.it.bi * p.v_descale_batch_stride + g.hk * p.v_descale_head_stride] : 1.f;
      // the row sum is the sum of the halves' sums (same stale max in both, so they add without a rescale); a + b == b + a
      // exactly, so both threads normalise by the same number
      float l_tot = l_run + pair_swap(l_run);
      const bool dead_row = (l_tot == 0.f) || (l_tot != l_tot);
      const float inv = dead_row ? 0.f : out_scale / l_tot;
      if (FP8) l_tot *= exp2f(-kFp8Offset);
      const float lse_val = dead_row ? -INFINITY : (m_run * scale_log2 + log2f(l_tot)) * kLn2;
      const int64_t row_idx = g.seq.q_off + pos;
      float w_prev = 0.f, w_cur = inv;
      const bool merging = split && p.merge != 0;
      const bool to_out = merging && p.merge == 2;
      const int64_t li = g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx;
      float lse_new = -INFINITY;
      if (merging && row_valid) {
        const float lse_prev = p.lse_accum[li];
        const float mx = fmaxf(lse_prev, lse_val);
        if (mx != -INFINITY) {
          const float e_prev = __expf(lse_prev - mx), e_cur = __expf(lse_val - mx);
          lse_new = mx + __logf(e_prev + e_cur);
          w_prev = __expf(lse_prev - lse_new);
          w_cur = __expf(lse_val - lse_new) * inv;
        } else w_cur = 0.f;
      }
      // when merging in place the peer must have read the stored LSE before half 0 replaces it
      if (merging) named_bar_sync(pair_bar, 64);
      if (half == 0 && row_valid) {
        if (merging) { if (to_out) p.lse[li] = lse_new; else p.lse_accum[li] = lse_new; }
        else if (!split) p.lse[li] = lse_val;
        else p.lse_accum[part * p.lseacc_split_stride + li] = lse_val;
      }
      if (live) { mbar_wait(&bar->o_full[t], o_it & 1, 0x80 + t); ++o_it; tc_fence_after(); }
      uint4* o16 = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.o) + g.seq.bq * p.o_batch_stride +
                                            row_idx * p.o_row_stride + hq * p.o_head_stride);
      float4* o32 = reinterpret_cast<float4*>(p.o_accum + part * p.oacc_split_stride + g.it.bi * p.oacc_batch_stride +
                                              hq * p.oacc_head_stride + row_idx * p.dv);
      if (HD <= 128 && p.o_tma) {
        // per 64-column slice the two threads of a row convert 32 columns each into the tile's [128][128 B] staging box
        uint8_t* stage = sO + t * FwdSmem<HD, ELT, false>::kOStageBytes;
        const bool issuer = (half == 0 && q == 0 && lane == 0);
#pragma unroll
        for (int sl = 0; sl < HD / 64; ++sl) {
          if (sl * 64 >= p.dv) break;
          uint32_t pk[16];
          {
            uint32_t o[32];
            if (live) { tmem_ld32(tO + sl * 64 + half * 32, o); tmem_ld_wait(); }
            else {
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = 0u;
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float a = __uint_as_float(o[2 * i]) * inv, b2 = __uint_as_float(o[2 * i + 1]) * inv;
              pk[i] = F16 ? pack_f16(a, b2) : pack_bf16(a, b2);
            }
          }
          if (issuer) tma_store_wait_read0();
          named_bar_sync(1 + t, 256);
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            uint4 u = make_uint4(pk[ch * 4], pk[ch * 4 + 1], pk[ch * 4 + 2], pk[ch * 4 + 3]);
            *reinterpret_cast<uint4*>(stage + row * 128 + (((half * 4 + ch) ^ (row & 7)) << 4)) = u;
          }
          fence_proxy_async_smem();
          named_bar_sync(1 + t, 256);
          if (issuer) {
            tma_store_4d(tmO, stage, sl * 64, pos - in_tile, hq, g.seq.bq);
            tma_store_commit();
          }
        }
      } else
#pragma unroll
      for (int cc = 0; cc < kOChunks; ++cc) {
        const int c = c0 + cc;
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
        if (lane == 0) mbar_arrive(&bar->o_empty[t]);
      }
    }
  }
  if (HD <= 128 && p.o_tma && half == 0 && q == 0 && lane == 0) tma_store_wait0();
  FA_PROF(7);
  FA_PROF_FLUSH();
}

// ----------------------------------------------------------------------------- kernel entry
// Register pool: 640 threads x 96 at launch; the issuing / producer warpgroup drops to kSplitOtherRegs and the sixteen
// softmax warps rise to kSplitSoftmaxRegs (512 x 112 + 128 x 32 = 61440).
#ifndef FA100_SPLIT_SMX_REGS
#define FA100_SPLIT_SMX_REGS 112
#define FA100_SPLIT_OTHER_REGS 32
#endif
constexpr int kSplitRegs[2] = {FA100_SPLIT_SMX_REGS, FA100_SPLIT_OTHER_REGS};

template <int HD, int ELT, bool FP8, bool F16>
__global__ void __launch_bounds__(HD > 128 ? kThreads : kSplitThreads, 1)
fa100_fwd_split_kernel(const __grid_constant__ FwdKernelParams p, const __grid_constant__ CUtensorMap tmQ,
                       const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                       const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmQv, uint32_t fmt) {
  using L = FwdSmem<HD, ELT, false>;
  constexpr bool SOLO = HD > 128;  // one Q tile per work item: 8 softmax warps (the default kernel's 12-warp CTA), else 16 (20 warps)
  constexpr int kSmxWarps = SOLO ? kSoftmaxWarps : kSplitSoftmaxWarps;
  constexpr int kMma = kSmxWarps, kLoad = kSmxWarps + 1, kPv = kSmxWarps + 2;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + L::kQSlots * L::kQBytes;
  uint8_t* sO = sKV + L::kStages * L::kKVBytes;
  float* xch = reinterpret_cast<float*>(smem + L::kDataBytes);
  FwdBarriers* bar = reinterpret_cast<FwdBarriers*>(smem + L::kDataBytes + FwdSmemSplit<HD, ELT>::kXchBytes);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  FA_TRACE_BEGIN();
  if (smem_u32(smem) & 1023u) __trap();
  if (warp == kLoad && lane == 0) { tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); if (p.o_tma) tma_prefetch_desc(&tmO); }
  if (warp == kMma && lane == 0) {
    constexpr int kSmxArrivals = 8;  // eight softmax warps per Q tile
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      mbar_init(&bar->q_full[t], 1); mbar_init(&bar->q_empty[t], 1);
      mbar_init(&bar->s_full[t], 1); mbar_init(&bar->p_full[t], kSmxArrivals); mbar_init(&bar->pv_done[t], 1);
      mbar_init(&bar->o_full[t], 1); mbar_init(&bar->o_empty[t], kSmxArrivals);
    }
    mbar_init(&bar->s_free, kSmxArrivals);
    for (int s = 0; s < 8; ++s) { mbar_init(&bar->kv_full[s], 1); mbar_init(&bar->kv_empty[s], 1); }
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc(&bar->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  pdl_wait();
  pdl_launch_dependents();
  if (p.sched && !sched_matches(p)) {
    if (threadIdx.x == 0 && blockIdx.x == 0) { g_fa100_hang_info[0] = 0xDEAD0000u | 0x5CEDu; g_fa100_hang_info[1] = 0; g_fa100_hang_info[2] = 0; g_fa100_hang_info[3] = 0; __threadfence(); }
    __trap();
  }
  if (warp < kSmxWarps) {
    setmaxnreg_inc<SOLO ? 224 : kSplitRegs[0]>();
    softmax_split_role<HD, ELT, FP8, F16, SOLO>(p, &tmO, sO, xch, bar, tmem, warp, lane);
  } else {
    setmaxnreg_dec<SOLO ? 56 : kSplitRegs[1]>();
    if (warp == kMma) mma_role<HD, ELT, FP8, false, true, false>(p, fmt, sQ, sKV, bar, tmem);
    if (warp == kPv) mma_role<HD, ELT, FP8, false, false, false>(p, fmt, sQ, sKV, bar, tmem);
    if (warp == kLoad) producer_role<HD, ELT, false, false>(p, &tmQ, &tmK, &tmV, sQ, sKV, bar, 0, &tmQv);
  }
  FA_TRACE_END();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace fa100
