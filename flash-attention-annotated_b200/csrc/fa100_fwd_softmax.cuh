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

// Online softmax follows hopper/softmax.h:92-168 (exp2 with the scale folded in, fp32 running max / sum,
// LSE = max*scale + ln(sum)) with Blackwell's row-per-lane TMEM layout and lazy rescaling.  Softmax warp w (0..15):
// Q tile t = w / 8, key-column half h = (w / 4) % 2, lane quarter q = w % 4 -> rows 32q.. of tile t, S columns [64h, +64).
template <int HD, int ELT, bool FP8, bool F16>
FA_DEVICE void softmax_role(const FwdKernelParams& p, FwdBarriers* bar, uint32_t tmem, int warp, int lane) {
  const int t = warp >> 3, h = (warp >> 2) & 1, q = warp & 3;
  const int row = q * 32 + lane;
  const uint32_t lane_base = uint32_t(q * 32) << 16;
  constexpr int kOCols = HD / 2;  // O columns this thread rescales / stores
  const uint32_t tS = tmem + lane_base + t * 128;
  const uint32_t tO = tmem + lane_base + 256 + t * 128 + h * kOCols;
  uint32_t s_it = 0, o_it = 0, xstep = 0;

  // hand a float to the partner thread of this row (same tile and row, other column half) and get theirs
  auto exchange = [&](float mine) -> float {
    float* slot = &bar->xchg[t][xstep & 1][0][0];
    slot[h * 128 + row] = mine;
    named_bar_sync(1 + t * 4 + q, 64);
    const float theirs = slot[(1 - h) * 128 + row];
    ++xstep;
    return theirs;
  };

  for (int w = blockIdx.x; w < p.num_work; w += gridDim.x) {
    const WorkGeom g = make_geom(p, w);
    const TileRange tr = g.tr[t];
    if (!tr.has_rows) continue;
    const int pos0 = (g.it.m_blk * kQTiles + t) * p.pos_per_tile;
    const int in_tile = p.pack > 1 ? row / p.pack : row;
    const int pos = pos0 + in_tile;
    const int hq = p.pack > 1 ? g.it.head * p.qheads_per_kv + row % p.pack : g.it.head;
    const bool row_valid = in_tile < p.pos_per_tile && pos < g.seq.sq;
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
    float scale_log2 = p.scale_log2, out_scale = 1.0f;
    if (FP8) {
      const float qd = p.q_descale ? p.q_descale[g.it.bi * p.q_descale_batch_stride + g.hk * p.q_descale_head_stride] : 1.f;
      const float kd = p.k_descale ? p.k_descale[g.it.bi * p.k_descale_batch_stride + g.hk * p.k_descale_head_stride] : 1.f;
      out_scale = p.v_descale ? p.v_descale[g.it.bi * p.v_descale_batch_stride + g.hk * p.v_descale_head_stride] : 1.f;
      scale_log2 *= qd * kd;  // softcap + fp8 is rejected on the host
    }

    float m_run = -INFINITY, l_run = 0.f;
    for (int nb = tr.n_max - 1; nb >= tr.n_min; --nb) {
      {
        mbar_wait(&bar->s_full[t], s_it & 1, 0x70 + t);
        ++s_it;
        tc_fence_after();
        // Thread h of a row owns the 32-column S chunks {h, h + 2}.  P (16-bit) overlays S columns [0, 64), i.e. the
        // LOW chunks of both threads, so the low chunk stays in registers across the row-max exchange (after that
        // barrier either thread may overwrite it); the high chunk (columns >= 64, never overwritten) is read twice,
        // once for the max and once for the exp.  Peak: ~48 live values per thread instead of 64 + 32.
        const int n_lo_chunk = nb * kBlockN + h * 32, n_hi_chunk = n_lo_chunk + 64;
        auto fixup = [&](uint32_t (&v)[32], int nbase) {  // softcap, then mask (hopper/mainloop...:1073-1077 order)
          if (p.softcap_pre != 0.f) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(fast_tanh(__uint_as_float(v[i]) * p.softcap_pre));
          }
          if (nbase + 32 > col_hi || nbase < col_lo) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (nbase + i >= col_hi || nbase + i < col_lo) v[i] = 0xff800000u;  // -inf
          }
        };
        auto max32 = [&](const uint32_t (&v)[32]) {
          float a = -INFINITY, b2 = -INFINITY;
#pragma unroll
          for (int i = 0; i < 32; i += 2) { a = fmaxf(a, __uint_as_float(v[i])); b2 = fmaxf(b2, __uint_as_float(v[i + 1])); }
          return fmaxf(a, b2);
        };
        uint32_t s_low[32];
        float m_half;
        {
          uint32_t s_high[32];
          tmem_ld32(tS + 64 + h * 32, s_high);
          tmem_ld32(tS + h * 32, s_low);
          tmem_ld_wait();
          fixup(s_high, n_hi_chunk);
          m_half = max32(s_high);
          fixup(s_low, n_lo_chunk);
          m_half = fmaxf(m_half, max32(s_low));
        }
        const float m_blk = fmaxf(m_half, exchange(m_half));
        FA_TRACE_POINT(0x90 + t);  // row max known
        const float m_cand = fmaxf(m_run, m_blk);
        // lazy rescale: keep the stale max unless it grew by more than 2^8 (NaN from -inf - -inf compares false)
        const bool grow = (m_cand - m_run) * scale_log2 > (FP8 ? 0.0f : kRescaleThreshold);
        const float acc_scale = grow ? exp2f((m_run - m_cand) * scale_log2) : 1.0f;
        if (grow) m_run = m_cand;
        const bool first = nb == tr.n_max - 1;
        if (!first && __any_sync(0xffffffffu, grow)) {
#pragma unroll
          for (int c = 0; c < kOCols / 32; ++c) {
            uint32_t o[32];
            tmem_ld32(tO + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * acc_scale);
            tmem_st32(tO + c * 32, o);
          }
        }
        const float m_sub = (m_run == -INFINITY) ? 0.f : m_run * scale_log2 - (FP8 ? 8.0f : 0.0f);
        // P = exp2(S * scale - max), packed to the operand dtype and written over S (columns [32h, +32) of the tile
        // for 16-bit operands, [16h, +16) for e4m3)
        float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
        auto exp_pack_store = [&](uint32_t (&sc)[32], int chunk) {  // chunk = index of the 32-key group inside the tile
#pragma unroll
          for (int i = 0; i < 32; ++i)
            sc[i] = __float_as_uint(ex2_approx(fmaf(__uint_as_float(sc[i]), scale_log2, -m_sub)));
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            sum0 += __uint_as_float(sc[i]); sum1 += __uint_as_float(sc[i + 1]);
            sum2 += __uint_as_float(sc[i + 2]); sum3 += __uint_as_float(sc[i + 3]);
          }
          if (!FP8) {
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float p0 = __uint_as_float(sc[2 * i]), p1 = __uint_as_float(sc[2 * i + 1]);
              pk[i] = F16 ? pack_f16(p0, p1) : pack_bf16(p0, p1);
            }
            tmem_st16(tS + chunk * 16, pk);
          } else {
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              uint16_t lo, hi;
              asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(__uint_as_float(sc[4 * i + 1])), "f"(__uint_as_float(sc[4 * i])));
              asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(__uint_as_float(sc[4 * i + 3])), "f"(__uint_as_float(sc[4 * i + 2])));
              pk[i] = uint32_t(lo) | (uint32_t(hi) << 16);
            }
            tmem_st8(tS + chunk * 8, pk);
          }
        };
        {
          uint32_t s_high[32];
          tmem_ld32(tS + 64 + h * 32, s_high);  // in flight while the low chunk is processed
          exp_pack_store(s_low, h);
          tmem_ld_wait();
          fixup(s_high, n_hi_chunk);
          exp_pack_store(s_high, 2 + h);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->p_full[t]);
        FA_TRACE_POINT(0xA0 + t);  // P published
        l_run = l_run * acc_scale + ((sum0 + sum1) + (sum2 + sum3));
      }
    }

    // ---- epilogue: O / rowsum, LSE (hopper/softmax.h:137-155, epilogue_fwd.hpp:215-481)
    const bool split = p.num_splits > 1;
    {
      const bool live = tr.n_max > tr.n_min;
      const float l_other = exchange(l_run);
      float l_tot = h == 0 ? l_run + l_other : l_other + l_run;  // same order in both threads
      const bool dead_row = (l_tot == 0.f) || (l_tot != l_tot);
      const float inv = dead_row ? 0.f : out_scale / l_tot;  // O carries the same 2^8 offset as the row sum (fp8)
      if (FP8) l_tot *= 0.00390625f;                          // undo the offset for the LSE
      const float lse_val = dead_row ? -INFINITY : (m_run * scale_log2 + log2f(l_tot)) * kLn2;
      const int64_t row_idx = g.seq.q_off + pos;
      if (row_valid && h == 0) {
        if (!split) p.lse[g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = lse_val;
        else p.lse_accum[g.it.split * p.lseacc_split_stride + g.it.bi * p.lse_batch_stride + hq * p.lse_head_stride + row_idx] = lse_val;
      }
      if (live) { mbar_wait(&bar->o_full[t], o_it & 1, 0x80 + t); ++o_it; tc_fence_after(); }
      uint4* o16 = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.o) + g.seq.bq * p.o_batch_stride +
                                            row_idx * p.o_row_stride + hq * p.o_head_stride + h * kOCols);
      float4* o32 = reinterpret_cast<float4*>(p.o_accum + g.it.split * p.oacc_split_stride + g.it.bi * p.oacc_batch_stride +
                                              hq * p.oacc_head_stride + row_idx * p.dv + h * kOCols);
#pragma unroll
      for (int c = 0; c < kOCols / 32; ++c) {
        uint32_t o[32];
        if (live) { tmem_ld32(tO + c * 32, o); tmem_ld_wait(); }
        else {
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = 0u;
        }
        if (row_valid) {
          if (!split) {
#pragma unroll
            for (int v8 = 0; v8 < 4; ++v8) {
              if (h * kOCols + c * 32 + v8 * 8 < p.dv) {
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
              if (h * kOCols + c * 32 + v4 * 4 < p.dv)
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
}

// ----------------------------------------------------------------------------- kernel entry
template <int HD, int ELT, bool FP8, bool F16>
__global__ void __launch_bounds__(kThreads, 1)
fa100_fwd_kernel(const __grid_constant__ FwdKernelParams p, const __grid_constant__ CUtensorMap tmQ,
                 const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, uint32_t fmt) {
  using L = FwdSmem<HD, ELT>;
  extern __shared__ __align__(1024) uint8_t smem[];  // 128-byte swizzle atoms need 1 KB alignment
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + kQTiles * L::kQBytes;
  FwdBarriers* bar = reinterpret_cast<FwdBarriers*>(smem + L::kDataBytes);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  FA_TRACE_BEGIN();
  if (smem_u32(smem) & 1023u) __trap();

  if (warp == kLoadWarp && lane == 0) { tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); }
  if (warp == kMmaWarp && lane == 0) {
#pragma unroll
    for (int t = 0; t < kQTiles; ++t) {
      mbar_init(&bar->q_full[t], 1); mbar_init(&bar->q_empty[t], 1);
      mbar_init(&bar->s_full[t], 1); mbar_init(&bar->p_full[t], 8);
      mbar_init(&bar->o_full[t], 1); mbar_init(&bar->o_empty[t], 8);
    }
    for (int s = 0; s < 8; ++s) { mbar_init(&bar->kv_full[s], 1); mbar_init(&bar->kv_empty[s], kQTiles); }
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc(&bar->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  // register pool at launch: 640 threads x 96; warps 16-19 give 32 each to the 16 softmax warps (+8 each)
  if (warp < kSoftmaxWarps) {
    setmaxnreg_inc<104>();
    softmax_role<HD, ELT, FP8, F16>(p, bar, tmem, warp, lane);
  } else {
    setmaxnreg_dec<64>();
    if (warp == kMmaWarp || warp == kMmaWarp + 1) mma_role<HD, ELT, FP8>(p, fmt, warp - kMmaWarp, sQ, sKV, bar, tmem);
    if (warp == kLoadWarp) producer_role<HD, ELT>(p, &tmQ, &tmK, &tmV, sQ, sKV, bar);
  }
  FA_TRACE_END();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace fa100
