// fa100_bwd.cu -- backward pass of the attention forward.
//
// Role of hopper/flash_bwd_launch_template.h + mainloop_bwd_sm90_tma_gmma_ws.hpp + flash_bwd_preprocess_kernel.h (and of
// flash_attn/cute/flash_bwd_sm100.py) re-designed around tensor memory, with the same math:
//   D_i  = sum_j dO_ij O_ij                      (preprocess; "softmax_d")
//   P    = exp(S * scale - LSE),  S = Q K^T       (recomputed, never stored)
//   dP   = dO V^T,   dS = P o (dP - D)
//   dV   = P^T dO,   dK = scale * dS^T Q,   dQ = scale * dS K
// Three launches, no atomics, bitwise deterministic:
//   fa100_bwd_dot_*kernel    D, memory-bound (16-byte loads, two rows per warp pass; a scalar fallback for unaligned rows)
//   fa100_bwd_dq_kernel      one CTA per (128-row query block, head, batch) walks the key blocks: S and dP in TMEM (fp32), the
//                            workers turn them into dS (16 bit, in its own double-buffered TMEM columns), dQ += dS K with dS as
//                            the TMEM A operand -- the forward's P V form; the next block's S / dP overlap this block's dS
//   fa100_bwd_dkdv_kernel    one CTA per (128-key block, KV head, batch) walks the query blocks of every query head of the
//                            group: S^T = K Q^T and dP^T = V dO^T (keys on the TMEM lanes), the workers make P^T, then dS^T
//                            (two phases pipelined against the tensor pipe), dV += P^T dO and dK += dS^T Q, again in the
//                            forward's P V form; the GQA sum over the group's heads happens in the accumulators
// Recomputing S and dP in both kernels costs 7 instead of 5 MMAs per (query block, key block) pair and buys: no dQ atomics,
// no transposes, only operand forms the forward kernel already uses (SS K-major x K-major, TS with an MN-major B).
// Both main kernels have the forward's role structure: 16 worker warps (four threads per TMEM lane), a TMA warp, an MMA-issuing
// warp, mbarrier hand-offs (DESIGN.md 4.8 has the pipelines and what the jitter build found in them).
#include "fa100_host.h"
#include "fa100_ptx.cuh"

#include <algorithm>
#include <atomic>
#include <cstring>

namespace fa100 {

constexpr float kLog2e = 1.4426950408889634f;
#ifndef FA100_BWD_PARTS
#define FA100_BWD_PARTS 4
#endif
constexpr int kBwdParts = FA100_BWD_PARTS;                // worker threads per TMEM lane (row): 2 or 4
constexpr int kBwdCols = 128 / kBwdParts;                 // columns of a step each of them owns
constexpr int kBwdC = kBwdCols / 32;                      // ... in 32-column register chunks
constexpr int kBwdWorkerWarps = 4 * kBwdParts;
constexpr int kBwdProducerWarp = kBwdWorkerWarps, kBwdMmaWarp = kBwdWorkerWarps + 1;
constexpr int kBwdRowSlots = 4;                           // dK/dV kernel: steps whose per-row constants sit in shared memory
constexpr int kBwdThreads = (kBwdWorkerWarps + 2) * 32;
constexpr int kBwdTile = 128;                 // query rows / keys per tile
constexpr int kBwdTileBytes = 128 * 128 * 2;  // one [128][128] 16-bit tile (two 64-column boxes)
constexpr int kBwdBoxBytes = 128 * 128;       // one [128 rows][128 B] box

struct BwdKernelParams {
  void* dq; void* dk; void* dv;
  const float* lse; const float* dsum;  // (b, h, seqlen_q) and (b, h, seqlen_q_rounded)
  int64_t dq_batch_stride, dq_row_stride, dq_head_stride;
  int64_t dk_batch_stride, dk_row_stride, dk_head_stride;
  int64_t dv_batch_stride, dv_row_stride, dv_head_stride;
  int32_t b, h, h_k, seqlen_q, seqlen_k, d;
  int32_t boxes;       // 64-column boxes per tile actually loaded: ceil(d / 64)
  int32_t slices;      // MMA k-slices over the head dim: ceil(d / 16)
  int32_t n16;         // d rounded up to 16: N of the products whose result has d columns
  int32_t f16;
  int32_t window_left, window_right;   // keys [i + shift - window_left, i + shift + window_right] are visible; < 0 = unbounded (causal: right = 0)
  float scale, scale_log2;
  float softcap_pre, softcap_log2;     // softcap: scale / softcap and softcap * log2(e); 0 = off
  const int32_t* cu_seqlens_q; const int32_t* cu_seqlens_k; const int32_t* seqused_q; const int32_t* seqused_k;
  int64_t lse_batch_stride, lse_head_stride, dsum_batch_stride, dsum_head_stride;  // fp32 elements; row stride 1
};

// per-batch lengths and offsets (hopper/seqlen.h:13-29); packed tensors use batch coordinate 0 and a row offset
struct BwdSeq { int sq, sk, q_off, k_off, bq, bk; };
FA_DEVICE BwdSeq bwd_seq(const BwdKernelParams& p, int bi) {
  BwdSeq s;
  s.q_off = p.cu_seqlens_q ? p.cu_seqlens_q[bi] : 0;
  s.k_off = p.cu_seqlens_k ? p.cu_seqlens_k[bi] : 0;
  s.sq = p.seqused_q ? p.seqused_q[bi] : (p.cu_seqlens_q ? p.cu_seqlens_q[bi + 1] - s.q_off : p.seqlen_q);
  s.sk = p.seqused_k ? p.seqused_k[bi] : (p.cu_seqlens_k ? p.cu_seqlens_k[bi + 1] - s.k_off : p.seqlen_k);
  s.bq = p.cu_seqlens_q ? 0 : bi;
  s.bk = p.cu_seqlens_k ? 0 : bi;
  return s;
}
// P and the softcap derivative factor for one score: P = exp(score - LSE), score = s * scale or softcap * tanh(s * scale / softcap)
FA_DEVICE float bwd_prob(const BwdKernelParams& p, float s, float lse2, float& dfac) {
  if (p.softcap_pre != 0.f) {
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(s * p.softcap_pre));
    dfac = 1.f - t * t;
    return exp2f(t * p.softcap_log2 - lse2);
  }
  dfac = 1.f;
  return exp2f(s * p.scale_log2 - lse2);
}

struct BwdBarriers {
  uint64_t fixed_full;        // the tile(s) loaded once per CTA
  uint64_t full[3], empty[3]; // ring of streamed tiles (dQ kernel: three K stages; dK/dV kernel: two Q + dO stages)
  uint64_t v_full[2], v_empty[2];  // dQ kernel: the V ring
  uint64_t s_ready;           // S / dP of a step are in tensor memory (tcgen05.commit)
  uint64_t regs_loaded;       // every worker warp holds the step's S / dP in registers (dQ kernel)
  uint64_t ops_ready[2];      // every worker warp has stored the step's 16-bit operands (dS; P^T and dS^T).  Two of them in the dQ
                              // kernel (block n uses n & 1): the workers may finish block n + 1 before the MMA warp has looked at block n
  uint64_t mma_done;          // accumulators final
  uint64_t dp_ready, p_ready; // dK/dV kernel: dP^T in tensor memory (commit); P^T stored by every worker warp
  uint64_t rows_full[4], rows_empty[4];  // dK/dV kernel: the step's per-query-row constants in shared memory
  uint32_t tmem_base;
  uint32_t pad_[3];
};

static_assert(sizeof(BwdBarriers) <= 256, "barrier block outgrew its slot");

template <bool F16> FA_DEVICE uint32_t bwd_pack(float a, float b) {
  if (F16) { __half2 v = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&v); }
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
// packed f32x2 arithmetic for the lean paths (one issue slot for two columns)
FA_DEVICE void bwd_fma2(float& a, float& b, float sc, float c0, float c1) {   // (a, b) = (a, b) * sc + (c0, c1)
  asm("{\n\t.reg .b64 x, s2, c2;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 s2, {%2, %2};\n\tmov.b64 c2, {%3, %4};\n\t"
      "fma.rn.ftz.f32x2 x, x, s2, c2;\n\tmov.b64 {%0, %1}, x;\n\t}" : "+f"(a), "+f"(b) : "f"(sc), "f"(c0), "f"(c1));
}
FA_DEVICE void bwd_dsv2(float& o0, float& o1, float p0, float p1, float dp0, float dp1, float d0, float d1) {  // (o0, o1) = (p0, p1) * ((dp0, dp1) - (d0, d1))
  asm("{\n\t.reg .b64 x, y, z;\n\tmov.b64 x, {%2, %3};\n\tmov.b64 y, {%4, %5};\n\tmov.b64 z, {%6, %7};\n\t"
      "sub.rn.ftz.f32x2 y, y, z;\n\tmul.rn.ftz.f32x2 x, x, y;\n\tmov.b64 {%0, %1}, x;\n\t}"
      : "=f"(o0), "=f"(o1) : "f"(p0), "f"(p1), "f"(dp0), "f"(dp1), "f"(d0), "f"(d1));
}

// ----------------------------------------------------------------------------- preprocess: D = rowsum(dO o O)
template <typename T>
__global__ void __launch_bounds__(256) fa100_bwd_dot_kernel(const T* __restrict__ dout, const T* __restrict__ out, float* __restrict__ dsum,
                                                            int64_t do_b, int64_t do_r, int64_t do_h, int64_t o_b, int64_t o_r, int64_t o_h,
                                                            int b, int h, int sq, int sq_rounded, int d) {
  asm volatile("griddepcontrol.wait;\n" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  const int lane = threadIdx.x & 31;
  const int64_t rows = int64_t(b) * h * sq_rounded;
  for (int64_t r = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; r < rows; r += (int64_t(gridDim.x) * blockDim.x) >> 5) {
    const int i = int(r % sq_rounded), hh = int((r / sq_rounded) % h), bb = int(r / (int64_t(sq_rounded) * h));
    float acc = 0.f;
    if (i < sq) {
      const T* pd = dout + bb * do_b + int64_t(i) * do_r + hh * do_h;
      const T* po = out + bb * o_b + int64_t(i) * o_r + hh * o_h;
      for (int c = lane; c < d; c += 32) acc += float(pd[c]) * float(po[c]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) dsum[r] = acc;  // rows past seqlen_q read as 0
  }
}

// the same with 16-byte loads: 16 lanes per row (8 elements each), two rows per warp pass.  Needs 16-byte aligned rows of dO and O
// (the host checks); 1.07 GB read at C2: 0.55 ms with the kernel above, memory-bound here
template <typename T>
__global__ void __launch_bounds__(256) fa100_bwd_dot_vec_kernel(const T* __restrict__ dout, const T* __restrict__ out, float* __restrict__ dsum,
                                                                int64_t do_b, int64_t do_r, int64_t do_h, int64_t o_b, int64_t o_r, int64_t o_h,
                                                                int b, int h, int sq, int sq_rounded, int d) {
  asm volatile("griddepcontrol.wait;\n" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  const int lane = threadIdx.x & 31, sub = lane & 15, half = lane >> 4;
  const int64_t rows = int64_t(b) * h * sq_rounded;
  const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5, nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r2 = warp0 * 2; r2 < rows; r2 += nwarps * 2) {
    const int64_t r = r2 + half;
    float acc = 0.f;
    if (r < rows) {
      const int i = int(r % sq_rounded), hh = int((r / sq_rounded) % h), bb = int(r / (int64_t(sq_rounded) * h));
      if (i < sq && sub * 8 < d) {
        const uint4 a = *reinterpret_cast<const uint4*>(dout + bb * do_b + int64_t(i) * do_r + hh * do_h + sub * 8);
        const uint4 c = *reinterpret_cast<const uint4*>(out + bb * o_b + int64_t(i) * o_r + hh * o_h + sub * 8);
        const T* pa = reinterpret_cast<const T*>(&a);
        const T* pc = reinterpret_cast<const T*>(&c);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc += float(pa[e]) * float(pc[e]);
      }
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (sub == 0 && r < rows) dsum[r] = acc;  // rows past seqlen_q read as 0
  }
}

// one [128 rows][d] tile = `boxes` TMA boxes of 64 columns; rows / columns past the tensor come back as zeros
FA_DEVICE void bwd_load_tile(uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int boxes, int row0, int head, int batch) {
  for (int hb = 0; hb < boxes; ++hb) tma_load_4d(dst + hb * kBwdBoxBytes, tm, bar, hb * 64, row0, head, batch);
}

// ----------------------------------------------------------------------------- roles shared by the two main kernels
// 8 worker warps (two threads per TMEM lane: warp w and w + 4 own the same 32 rows and 64 of a step's 128 columns each), a
// TMA warp and an MMA-issuing warp.  Hand-offs are mbarriers as in the forward kernel: s_ready (tcgen05.commit), regs_loaded
// and ops_ready (one arrive per worker warp), full / empty for the two shared-memory stages.
FA_DEVICE void bwd_warp_arrive(uint64_t* bar, int lane) {
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(bar);
}
// the 16-bit operand a worker leaves in tensor memory: kBwdCols / 2 columns
FA_DEVICE void bwd_tmem_st(uint32_t addr, const uint32_t (&r)[32]) { tmem_st32(addr, r); }
FA_DEVICE void bwd_tmem_st(uint32_t addr, const uint32_t (&r)[16]) { tmem_st16(addr, r); }
// TMEM column offset of MMA k-slice ks (16 consecutive columns of the step) inside a 128-column region whose owners wrote
// their kBwdCols / 2 operand columns over the start of their own kBwdCols columns
FA_DEVICE constexpr uint32_t bwd_slice_col(int ks) { return uint32_t((ks * 16 / kBwdCols) * kBwdCols + (ks * 16 % kBwdCols) / 2); }
FA_DEVICE float bwd_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// ----------------------------------------------------------------------------- dQ
// TMEM: S @0, dP @128 (fp32, 128 columns each), dQ @256, dS (16 bit, 64 columns) double-buffered @384 / @448.  With dS in its
// own columns S and dP are free as soon as the workers hold them in registers, so the MMA warp computes the next block's S
// and dP while the workers turn this block's into dS; the order in the tensor pipe is S,dP(n+1) | dQ(n) | S,dP(n+2) | ...
template <bool F16>
__global__ void __launch_bounds__(kBwdThreads, 1)
fa100_bwd_dq_kernel(const __grid_constant__ BwdKernelParams p, const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, uint32_t fmt) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;                          // Q tile
  uint8_t* sdO = smem + kBwdTileBytes;         // dO tile
  uint8_t* sK = smem + 2 * kBwdTileBytes;      // K stages 0, 1, 2: a K block is read twice, by S (early) and by dQ (one block later)
  uint8_t* sV = smem + 5 * kBwdTileBytes;      // V stages 0, 1: released as soon as dP is computed
  BwdBarriers* bar = reinterpret_cast<BwdBarriers*>(smem + 7 * kBwdTileBytes);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // the hardware hands out blockIdx.x in ascending order: under a causal mask the late row blocks are the long ones, so they go first
  const int m_blk = int(gridDim.x) - 1 - int(blockIdx.x), hq = blockIdx.y, bi = blockIdx.z;
  const int hk = hq / (p.h / p.h_k);
  if (tid == 0) {
    mbar_init(&bar->fixed_full, 1); mbar_init(&bar->s_ready, 1); mbar_init(&bar->mma_done, 1);
    mbar_init(&bar->regs_loaded, kBwdWorkerWarps); mbar_init(&bar->ops_ready[0], kBwdWorkerWarps); mbar_init(&bar->ops_ready[1], kBwdWorkerWarps);
    for (int s = 0; s < 3; ++s) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&bar->v_full[s], 1); mbar_init(&bar->v_empty[s], 1); }
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc(&bar->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  pdl_wait();
  pdl_launch_dependents();

  const BwdSeq sq_ = bwd_seq(p, bi);
  const int row0 = m_blk * kBwdTile;
  const int shift = sq_.sk - sq_.sq;
  // key blocks [n_min, n_max) any row of this tile can see (hopper/block.h:14-60 with 128-row tiles)
  int n_max = (sq_.sk + kBwdTile - 1) / kBwdTile, n_min = 0;
  if (p.window_right >= 0) n_max = min(n_max, max(0, (min(row0 + kBwdTile, sq_.sq) + shift + p.window_right + kBwdTile - 1) / kBwdTile));
  if (p.window_left >= 0) n_min = max(0, (row0 + shift - p.window_left) / kBwdTile);
  if (row0 >= sq_.sq) n_max = n_min;
  const int n_blocks = max(0, n_max - n_min);
  const uint32_t tile_tx = uint32_t(p.boxes) * kBwdBoxBytes;

  if (warp == kBwdProducerWarp) {
    // ---- TMA: Q and dO once, K / V blocks through the two stages
    if (lane == 0 && n_blocks > 0) {
      mbar_arrive_expect_tx(&bar->fixed_full, 2 * tile_tx);
      bwd_load_tile(sQ, &tmQ, &bar->fixed_full, p.boxes, sq_.q_off + row0, hq, sq_.bq);
      bwd_load_tile(sdO, &tmdO, &bar->fixed_full, p.boxes, sq_.q_off + row0, hq, sq_.bq);
      for (int nb = 0; nb < n_blocks; ++nb) {
        const int ks = nb % 3, vs = nb & 1;
        if (nb >= 3) mbar_wait(&bar->empty[ks], (nb / 3 - 1) & 1, 0xB1);
        mbar_arrive_expect_tx(&bar->full[ks], tile_tx);
        bwd_load_tile(sK + ks * kBwdTileBytes, &tmK, &bar->full[ks], p.boxes, sq_.k_off + (n_min + nb) * kBwdTile, hk, sq_.bk);
        if (nb >= 2) mbar_wait(&bar->v_empty[vs], (nb / 2 - 1) & 1, 0xB7);
        mbar_arrive_expect_tx(&bar->v_full[vs], tile_tx);
        bwd_load_tile(sV + vs * kBwdTileBytes, &tmV, &bar->v_full[vs], p.boxes, sq_.k_off + (n_min + nb) * kBwdTile, hk, sq_.bk);
      }
    }
  } else if (warp == kBwdMmaWarp) {
    // ---- MMA issue: the whole warp runs the control flow and the waits, one elected lane issues (descriptors stay in
    // uniform registers, no divergence code around the MMAs -- the forward kernel's idiom)
    if (n_blocks > 0) {
      const uint32_t idesc_s = make_idesc(fmt, fmt, kBwdTile, kBwdTile, 0, 0);   // S = Q K^T, dP = dO V^T
      const uint32_t idesc_dq = make_idesc(fmt, fmt, kBwdTile, p.n16, 0, 1);     // dQ += dS K (B = K tile, MN-major)
      constexpr uint32_t kHi = sdesc_hi(1024);
      constexpr uint32_t kLboK = sdesc_lo_lbo(16), kLboMN = sdesc_lo_lbo(kBwdBoxBytes);
      const uint32_t q_lo = (smem_u32(sQ) >> 4) | kLboK, do_lo = (smem_u32(sdO) >> 4) | kLboK;
      auto issue_s_dp = [&](int nb) {
        const int kst = nb % 3, vst = nb & 1;
        mbar_wait(&bar->full[kst], (nb / 3) & 1, 0xB2);
        mbar_wait(&bar->v_full[vst], (nb / 2) & 1, 0xB8);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t k_lo = (smem_u32(sK + kst * kBwdTileBytes) >> 4) | kLboK, v_lo = (smem_u32(sV + vst * kBwdTileBytes) >> 4) | kLboK;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t off = (ks / 4) * (kBwdBoxBytes >> 4) + (ks % 4) * 2;
            if (ks < p.slices) mma_ss<false>(tmem, q_lo + off, k_lo + off, kHi, idesc_s, ks > 0);
          }
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t off = (ks / 4) * (kBwdBoxBytes >> 4) + (ks % 4) * 2;
            if (ks < p.slices) mma_ss<false>(tmem + 128, do_lo + off, v_lo + off, kHi, idesc_s, ks > 0);
          }
          tc_commit(&bar->s_ready);
          tc_commit(&bar->v_empty[vst]);
        }
        __syncwarp();
      };
      mbar_wait(&bar->fixed_full, 0, 0xB0);
      issue_s_dp(0);
      for (int nb = 0; nb < n_blocks; ++nb) {
        const int st = nb & 1, kst = nb % 3;
        if (nb + 1 < n_blocks) {
          mbar_wait(&bar->regs_loaded, nb & 1, 0xB5);  // S / dP of block nb are in the workers' registers
          issue_s_dp(nb + 1);
        }
        mbar_wait(&bar->ops_ready[nb & 1], (nb >> 1) & 1, 0xB6);      // dS of block nb is in its buffer
        tc_fence_after();
        if (elect_one()) {
          const uint32_t k_mn = (smem_u32(sK + kst * kBwdTileBytes) >> 4) | kLboMN;
          const uint32_t t_ds = tmem + 384 + st * 64;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)  // contraction over the 128 keys, 16 per slice
            mma_ts<false>(tmem + 256, t_ds + ks * 8, k_mn + ks * (16 * 128 >> 4), kHi, idesc_dq, (nb > 0 || ks > 0));
          tc_commit(&bar->empty[kst]);
          if (nb == n_blocks - 1) tc_commit(&bar->mma_done);
        }
        __syncwarp();
      }
    }
  } else {
    // ---- workers: this thread's query row and its kBwdCols of a block's 128 key columns
    const int q4 = warp & 3, part = warp >> 2;
    const int i = row0 + q4 * 32 + lane;
    const bool row_ok = i < sq_.sq;
    float lse2 = INFINITY, dsum = 0.f;  // rows past the end and rows without keys (LSE = -inf) get P = 0
    if (row_ok) {
      const float l = p.lse[bi * p.lse_batch_stride + hq * p.lse_head_stride + sq_.q_off + i];
      lse2 = (l == -INFINITY) ? INFINITY : l * kLog2e;
      dsum = p.dsum[bi * p.dsum_batch_stride + hq * p.dsum_head_stride + sq_.q_off + i];
    }
    const int col_hi = p.window_right >= 0 ? min(sq_.sk, i + 1 + shift + p.window_right) : sq_.sk;
    const int col_lo = p.window_left >= 0 ? max(0, i + shift - p.window_left) : 0;
    const uint32_t tRow = tmem + (uint32_t(q4 * 32) << 16);
    const float neg_lse2 = -lse2;

    for (int nb = 0; nb < n_blocks; ++nb) {
      mbar_wait(&bar->s_ready, nb & 1, 0xB3);
      tc_fence_after();
      uint32_t s[kBwdC][32], dp[kBwdC][32], pk[kBwdCols / 2];
#pragma unroll
      for (int c = 0; c < kBwdC; ++c) { tmem_ld32(tRow + part * kBwdCols + c * 32, s[c]); tmem_ld32(tRow + 128 + part * kBwdCols + c * 32, dp[c]); }
      tmem_ld_wait();
      bwd_warp_arrive(&bar->regs_loaded, lane);
      const int n0 = (n_min + nb) * kBwdTile;
      // every row of the tile sees every key of the block and there is no softcap: the lean path (CTA-uniform test)
      const bool lean = p.softcap_pre == 0.f && n0 + kBwdTile <= sq_.sk &&
                        (p.window_right < 0 || n0 + kBwdTile <= row0 + 1 + shift + p.window_right) &&
                        (p.window_left < 0 || n0 >= row0 + kBwdTile - 1 + shift - p.window_left);
      if (lean) {
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float x0 = __uint_as_float(s[c][j]), x1 = __uint_as_float(s[c][j + 1]), o0, o1;
            bwd_fma2(x0, x1, p.scale_log2, neg_lse2, neg_lse2);
            bwd_dsv2(o0, o1, bwd_ex2(x0), bwd_ex2(x1), __uint_as_float(dp[c][j]), __uint_as_float(dp[c][j + 1]), dsum, dsum);
            pk[c * 16 + j / 2] = bwd_pack<F16>(o0, o1);
          }
      } else if (p.softcap_pre == 0.f) {
        // masked block without softcap (the diagonal block of every causal item): the lean arithmetic, then columns outside this
        // row's window [lo, hi) become 0 -- compile-time column index against two per-thread limits
        const int hi = col_hi - n0 - part * kBwdCols, lo = col_lo - n0 - part * kBwdCols;
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float x0 = __uint_as_float(s[c][j]), x1 = __uint_as_float(s[c][j + 1]), o0, o1;
            bwd_fma2(x0, x1, p.scale_log2, neg_lse2, neg_lse2);
            bwd_dsv2(o0, o1, bwd_ex2(x0), bwd_ex2(x1), __uint_as_float(dp[c][j]), __uint_as_float(dp[c][j + 1]), dsum, dsum);
            if (c * 32 + j >= hi || c * 32 + j < lo) o0 = 0.f;
            if (c * 32 + j + 1 >= hi || c * 32 + j + 1 < lo) o1 = 0.f;
            pk[c * 16 + j / 2] = bwd_pack<F16>(o0, o1);
          }
      } else {
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float ds[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int col = n0 + part * kBwdCols + c * 32 + j + e;
              float dfac = 0.f;
              const float pr = (col < col_hi && col >= col_lo) ? bwd_prob(p, __uint_as_float(s[c][j + e]), lse2, dfac) : 0.f;
              ds[e] = pr * (__uint_as_float(dp[c][j + e]) - dsum) * dfac;
            }
            pk[c * 16 + j / 2] = bwd_pack<F16>(ds[0], ds[1]);
          }
      }
      // keys [kBwdCols part, +kBwdCols) of the block are columns [kBwdCols / 2 part, ...) of the 16-bit dS buffer
      bwd_tmem_st(tRow + 384 + (nb & 1) * 64 + part * (kBwdCols / 2), pk);
      tmem_st_wait();
      bwd_warp_arrive(&bar->ops_ready[nb & 1], lane);
    }
    // dQ = scale * accumulator; each thread stores its share of the columns
    if (n_blocks > 0) { mbar_wait(&bar->mma_done, 0, 0xB4); tc_fence_after(); }
    uint16_t* dst = reinterpret_cast<uint16_t*>(p.dq) + sq_.bq * p.dq_batch_stride + int64_t(sq_.q_off + i) * p.dq_row_stride + hq * p.dq_head_stride;
    for (int c = part * kBwdC; c < (part + 1) * kBwdC; ++c) {
      if (c * 32 >= p.d) break;
      uint32_t o[32];
      if (n_blocks > 0) { tmem_ld32(tRow + 256 + c * 32, o); tmem_ld_wait(); }
      else {
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = 0u;
      }
      if (row_ok) {
#pragma unroll
        for (int v8 = 0; v8 < 4; ++v8)
          if (c * 32 + v8 * 8 < p.d) {
            uint4 u;
            uint32_t* up = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) up[e] = bwd_pack<F16>(__uint_as_float(o[v8 * 8 + 2 * e]) * p.scale, __uint_as_float(o[v8 * 8 + 2 * e + 1]) * p.scale);
            *reinterpret_cast<uint4*>(dst + c * 32 + v8 * 8) = u;
          }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// ----------------------------------------------------------------------------- dK, dV
// TMEM: A @0 and B @128 (128 fp32 columns each), dK @256, dV @384 -- all 512 columns.  S^T = K Q^T lands in A (keys on the lanes,
// the step's 128 query rows on the columns), dP^T = V dO^T in B; a worker turns its 64 columns of S^T into P^T (16 bit) and writes
// it over the first 32 of THOSE columns (nobody else reads them), likewise dS^T over dP^T, so no thread waits for another before it
// stores.  A step is split in two so that tensor work and arithmetic overlap:
//   tensor pipe:  ... dV(t) | S^T(t+1) | dK(t) | dP^T(t+1) | dV(t+1) | ...
//   workers:      ... P phase(t): exp  ->  dS phase(t): P o (dP - D)  ->  P phase(t+1) ...
// S^T(t+1) may overwrite A once dV(t) has read P^T(t) (in-order pipe), dP^T(t+1) may overwrite B once dK(t) has read dS^T(t).
// One CTA per (key block, KV head, batch); the steps run over (query head of the group, query block) and all accumulate into the
// same dK / dV, which is the GQA reduction.
template <bool F16>
__global__ void __launch_bounds__(kBwdThreads, 1)
fa100_bwd_dkdv_kernel(const __grid_constant__ BwdKernelParams p, const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                      const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, uint32_t fmt) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sK = smem;                          // K tile
  uint8_t* sV = smem + kBwdTileBytes;          // V tile
  uint8_t* sQ = smem + 2 * kBwdTileBytes;      // Q stages 0, 1
  uint8_t* sdO = smem + 4 * kBwdTileBytes;     // dO stages 0, 1
  float* sL = reinterpret_cast<float*>(smem + 6 * kBwdTileBytes);  // [4][128] -LSE * log2(e) of a step's query rows, fetched two steps ahead
  float* sD = sL + kBwdRowSlots * kBwdTile;                         // [4][128] D of the same rows
  BwdBarriers* bar = reinterpret_cast<BwdBarriers*>(smem + 6 * kBwdTileBytes + 2 * kBwdRowSlots * kBwdTile * 4);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_blk = blockIdx.x, hk = blockIdx.y, bi = blockIdx.z;
  const int G = p.h / p.h_k;
  if (tid == 0) {
    mbar_init(&bar->fixed_full, 1); mbar_init(&bar->s_ready, 1); mbar_init(&bar->dp_ready, 1); mbar_init(&bar->mma_done, 1);
    mbar_init(&bar->p_ready, kBwdWorkerWarps); mbar_init(&bar->ops_ready[0], kBwdWorkerWarps);
    for (int s_ = 0; s_ < 2; ++s_) {
      mbar_init(&bar->full[s_], 1); mbar_init(&bar->empty[s_], 1);       // Q stages
      mbar_init(&bar->v_full[s_], 1); mbar_init(&bar->v_empty[s_], 1);   // dO stages (free one product earlier than Q)
    }
    for (int s_ = 0; s_ < kBwdRowSlots; ++s_) { mbar_init(&bar->rows_full[s_], 1); mbar_init(&bar->rows_empty[s_], kBwdWorkerWarps); }
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc(&bar->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  pdl_wait();
  pdl_launch_dependents();

  const BwdSeq sq_ = bwd_seq(p, bi);
  const int n0 = n_blk * kBwdTile;
  const int shift = sq_.sk - sq_.sq;
  const int m_blocks = (sq_.sq + kBwdTile - 1) / kBwdTile;
  // query blocks [m_lo, m_hi) with a row that sees a key of this tile: row i sees keys [i + shift - wl, i + shift + wr]
  int m_lo = 0, m_hi = m_blocks;
  if (p.window_right >= 0) m_lo = min(m_blocks, max(0, (n0 - shift - p.window_right) / kBwdTile));
  if (p.window_left >= 0) {
    const int last_row = min(n0 + kBwdTile, sq_.sk) - 1 - shift + p.window_left;  // the last row that sees a key of this tile
    m_hi = last_row < 0 ? 0 : min(m_blocks, last_row / kBwdTile + 1);
  }
  if (n0 >= sq_.sk) m_hi = m_lo;
  const int per_head = max(0, m_hi - m_lo);
  const int steps = per_head * G;
  const uint32_t tile_tx = uint32_t(p.boxes) * kBwdBoxBytes;

  if (warp == kBwdProducerWarp) {
    // ---- TMA (one lane) and the per-row constants of every step (all lanes, four query rows each, two steps ahead)
    if (lane == 0 && steps > 0) {
      mbar_arrive_expect_tx(&bar->fixed_full, 2 * tile_tx);
      bwd_load_tile(sK, &tmK, &bar->fixed_full, p.boxes, sq_.k_off + n0, hk, sq_.bk);
      bwd_load_tile(sV, &tmV, &bar->fixed_full, p.boxes, sq_.k_off + n0, hk, sq_.bk);
    }
    auto fetch_rows = [&](int t) {
      const int slot = t % kBwdRowSlots;
      const int hq = hk * G + t / per_head, m = m_lo + t % per_head;
      if (t >= kBwdRowSlots) mbar_wait(&bar->rows_empty[slot], (t / kBwdRowSlots - 1) & 1, 0xC6);
      const float* lse = p.lse + bi * p.lse_batch_stride + hq * p.lse_head_stride + sq_.q_off;
      const float* dsm = p.dsum + bi * p.dsum_batch_stride + hq * p.dsum_head_stride + sq_.q_off;
      float l[4], d_[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {  // eight independent loads (clamped index: rows past the end are masked below)
        const int i = min(m * kBwdTile + r * 32 + lane, sq_.sq - 1);
        l[r] = lse[i]; d_[r] = dsm[i];
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const bool ok = m * kBwdTile + r * 32 + lane < sq_.sq;
        // rows past the end and rows without keys (LSE = -inf): P = exp2(-inf) = 0
        sL[slot * kBwdTile + r * 32 + lane] = (ok && l[r] != -INFINITY) ? -l[r] * kLog2e : -INFINITY;
        sD[slot * kBwdTile + r * 32 + lane] = ok ? d_[r] : 0.f;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->rows_full[slot]);
    };
    // the first two steps' tiles go out before anything else (no stage to wait for): fetching their row constants first cost every
    // CTA ~3 us before its first S^T could start -- a third of a 1024-token CTA's life
    auto load_step = [&](int t) {
      const int st = t & 1;
      const int hq = hk * G + t / per_head, m = m_lo + t % per_head;
      if (lane == 0) {
        if (t >= 2) mbar_wait(&bar->v_empty[st], (t / 2 - 1) & 1, 0xCA);
        mbar_arrive_expect_tx(&bar->v_full[st], tile_tx);
        bwd_load_tile(sdO + st * kBwdTileBytes, &tmdO, &bar->v_full[st], p.boxes, sq_.q_off + m * kBwdTile, hq, sq_.bq);
        if (t >= 2) mbar_wait(&bar->empty[st], (t / 2 - 1) & 1, 0xC1);
        mbar_arrive_expect_tx(&bar->full[st], tile_tx);
        bwd_load_tile(sQ + st * kBwdTileBytes, &tmQ, &bar->full[st], p.boxes, sq_.q_off + m * kBwdTile, hq, sq_.bq);
      }
      __syncwarp();
    };
    if (steps > 0) load_step(0);
    if (steps > 1) load_step(1);
    if (steps > 0) fetch_rows(0);
    if (steps > 1) fetch_rows(1);
    for (int t = 0; t < steps; ++t) {
      if (t >= 2) load_step(t);
      if (t + 2 < steps) fetch_rows(t + 2);
    }
  } else if (warp == kBwdMmaWarp) {
    if (steps > 0) {
      const uint32_t idesc_s = make_idesc(fmt, fmt, kBwdTile, kBwdTile, 0, 0);
      const uint32_t idesc_acc = make_idesc(fmt, fmt, kBwdTile, p.n16, 0, 1);
      constexpr uint32_t kHi = sdesc_hi(1024);
      constexpr uint32_t kLboK = sdesc_lo_lbo(16), kLboMN = sdesc_lo_lbo(kBwdBoxBytes);
      const uint32_t k_lo = (smem_u32(sK) >> 4) | kLboK, v_lo = (smem_u32(sV) >> 4) | kLboK;
      auto issue_s = [&](int t) {   // S^T(t) = K Q(t)^T -> A
        const int st = t & 1;
        mbar_wait(&bar->full[st], (t / 2) & 1, 0xC2);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t q_lo = (smem_u32(sQ + st * kBwdTileBytes) >> 4) | kLboK;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t off = (ks / 4) * (kBwdBoxBytes >> 4) + (ks % 4) * 2;
            if (ks < p.slices) mma_ss<false>(tmem, k_lo + off, q_lo + off, kHi, idesc_s, ks > 0);
          }
          tc_commit(&bar->s_ready);
        }
        __syncwarp();
      };
      auto issue_dp = [&](int t) {  // dP^T(t) = V dO(t)^T -> B
        const int st = t & 1;
        mbar_wait(&bar->v_full[st], (t / 2) & 1, 0xCB);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t do_lo = (smem_u32(sdO + st * kBwdTileBytes) >> 4) | kLboK;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t off = (ks / 4) * (kBwdBoxBytes >> 4) + (ks % 4) * 2;
            if (ks < p.slices) mma_ss<false>(tmem + 128, v_lo + off, do_lo + off, kHi, idesc_s, ks > 0);
          }
          tc_commit(&bar->dp_ready);
        }
        __syncwarp();
      };
      mbar_wait(&bar->fixed_full, 0, 0xC0);
      issue_s(0);
      issue_dp(0);
      for (int t = 0; t < steps; ++t) {
        const int st = t & 1;
        mbar_wait(&bar->p_ready, t & 1, 0xC7);      // P^T(t) is in A
        tc_fence_after();
        if (elect_one()) {
          const uint32_t do_mn = (smem_u32(sdO + st * kBwdTileBytes) >> 4) | kLboMN;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)  // contraction over the 128 query rows of the step (bwd_slice_col: where a slice's P^T sits)
            mma_ts<false>(tmem + 384, tmem + bwd_slice_col(ks), do_mn + ks * (16 * 128 >> 4), kHi, idesc_acc, (t > 0 || ks > 0));  // dV += P^T dO
          tc_commit(&bar->v_empty[st]);  // last reader of dO(t)
        }
        __syncwarp();
        if (t + 1 < steps) issue_s(t + 1);
        mbar_wait(&bar->ops_ready[0], t & 1, 0xC5);    // dS^T(t) is in B
        tc_fence_after();
        if (elect_one()) {
          const uint32_t q_mn = (smem_u32(sQ + st * kBwdTileBytes) >> 4) | kLboMN;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            mma_ts<false>(tmem + 256, tmem + 128 + bwd_slice_col(ks), q_mn + ks * (16 * 128 >> 4), kHi, idesc_acc, (t > 0 || ks > 0));  // dK += dS^T Q
          tc_commit(&bar->empty[st]);
          if (t == steps - 1) tc_commit(&bar->mma_done);
        }
        __syncwarp();
        if (t + 1 < steps) issue_dp(t + 1);
      }
    }
  } else {
    // ---- workers: this thread's key row and its kBwdCols of a step's 128 query columns
    const int q4 = warp & 3, part = warp >> 2;
    const int kj = n0 + q4 * 32 + lane;
    const uint32_t tA = tmem + (uint32_t(q4 * 32) << 16) + part * kBwdCols, tB = tA + 128;
    for (int t = 0; t < steps; ++t) {
      const int m = m_lo + t % per_head;
      const int slot = t % kBwdRowSlots;
      const float* nl = sL + slot * kBwdTile + part * kBwdCols;
      const float* dd = sD + slot * kBwdTile + part * kBwdCols;
      const int i0 = m * kBwdTile;
      // every key of the tile is visible from every query row of the step, no softcap: the lean path (CTA-uniform)
      const bool lean = p.softcap_pre == 0.f && (p.window_right < 0 || n0 + kBwdTile - 1 <= i0 + shift + p.window_right) &&
                        (p.window_left < 0 || n0 >= i0 + kBwdTile - 1 + shift - p.window_left);
      // query columns [q_lo, q_hi) of this thread's kBwdCols see its key row: row i sees keys [i + shift - wl, i + shift + wr]
      const int q_base = i0 + part * kBwdCols + shift;
      const int q_lo = p.window_right >= 0 ? kj - p.window_right - q_base : -(1 << 30);
      const int q_hi = p.window_left >= 0 ? kj + p.window_left - q_base + 1 : (1 << 30);
      const bool plain = lean || p.softcap_pre == 0.f;  // s holds P (masked where needed) after the P phase
      mbar_wait(&bar->rows_full[slot], (t / kBwdRowSlots) & 1, 0xC8);
      mbar_wait(&bar->s_ready, t & 1, 0xC3);
      tc_fence_after();
      uint32_t s[kBwdC][32], pk[kBwdCols / 2];
#pragma unroll
      for (int c = 0; c < kBwdC; ++c) tmem_ld32(tA + c * 32, s[c]);
      tmem_ld_wait();
      // ---- P phase: lean: s becomes P (fp32, kept for the dS phase); general: s stays raw and P is recomputed there
      if (lean) {
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 l4 = *reinterpret_cast<const float4*>(nl + c * 32 + j);
            float x0 = __uint_as_float(s[c][j]), x1 = __uint_as_float(s[c][j + 1]), x2 = __uint_as_float(s[c][j + 2]), x3 = __uint_as_float(s[c][j + 3]);
            bwd_fma2(x0, x1, p.scale_log2, l4.x, l4.y);
            bwd_fma2(x2, x3, p.scale_log2, l4.z, l4.w);
            const float p0 = bwd_ex2(x0), p1 = bwd_ex2(x1), p2 = bwd_ex2(x2), p3 = bwd_ex2(x3);
            s[c][j] = __float_as_uint(p0); s[c][j + 1] = __float_as_uint(p1); s[c][j + 2] = __float_as_uint(p2); s[c][j + 3] = __float_as_uint(p3);
            pk[c * 16 + j / 2] = bwd_pack<F16>(p0, p1);
            pk[c * 16 + j / 2 + 1] = bwd_pack<F16>(p2, p3);
          }
      } else if (p.softcap_pre == 0.f) {
        // masked step without softcap (the diagonal step of every causal key block): the lean arithmetic, then query columns that
        // do not see this key row become 0; s keeps the masked P for the dS phase
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 l4 = *reinterpret_cast<const float4*>(nl + c * 32 + j);
            float x0 = __uint_as_float(s[c][j]), x1 = __uint_as_float(s[c][j + 1]), x2 = __uint_as_float(s[c][j + 2]), x3 = __uint_as_float(s[c][j + 3]);
            bwd_fma2(x0, x1, p.scale_log2, l4.x, l4.y);
            bwd_fma2(x2, x3, p.scale_log2, l4.z, l4.w);
            float p0 = bwd_ex2(x0), p1 = bwd_ex2(x1), p2 = bwd_ex2(x2), p3 = bwd_ex2(x3);
            if (c * 32 + j >= q_hi || c * 32 + j < q_lo) p0 = 0.f;
            if (c * 32 + j + 1 >= q_hi || c * 32 + j + 1 < q_lo) p1 = 0.f;
            if (c * 32 + j + 2 >= q_hi || c * 32 + j + 2 < q_lo) p2 = 0.f;
            if (c * 32 + j + 3 >= q_hi || c * 32 + j + 3 < q_lo) p3 = 0.f;
            s[c][j] = __float_as_uint(p0); s[c][j + 1] = __float_as_uint(p1); s[c][j + 2] = __float_as_uint(p2); s[c][j + 3] = __float_as_uint(p3);
            pk[c * 16 + j / 2] = bwd_pack<F16>(p0, p1);
            pk[c * 16 + j / 2 + 1] = bwd_pack<F16>(p2, p3);
          }
      } else {
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float pr[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int col = c * 32 + j + e;
              const int pos = i0 + part * kBwdCols + col + shift;  // the key on this query row's diagonal
              const bool vis = (p.window_right < 0 || kj <= pos + p.window_right) && (p.window_left < 0 || kj >= pos - p.window_left);
              float dfac = 0.f;
              pr[e] = vis ? bwd_prob(p, __uint_as_float(s[c][j + e]), -nl[col], dfac) : 0.f;
            }
            pk[c * 16 + j / 2] = bwd_pack<F16>(pr[0], pr[1]);
          }
      }
      bwd_tmem_st(tA, pk);   // P^T over the first half of this thread's own S^T columns
      tmem_st_wait();
      bwd_warp_arrive(&bar->p_ready, lane);
      // ---- dS phase
      mbar_wait(&bar->dp_ready, t & 1, 0xC9);
      tc_fence_after();
      uint32_t dp[kBwdC][32];
#pragma unroll
      for (int c = 0; c < kBwdC; ++c) tmem_ld32(tB + c * 32, dp[c]);
      tmem_ld_wait();
      if (plain) {
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 d4 = *reinterpret_cast<const float4*>(dd + c * 32 + j);
            float o0, o1, o2, o3;
            bwd_dsv2(o0, o1, __uint_as_float(s[c][j]), __uint_as_float(s[c][j + 1]), __uint_as_float(dp[c][j]), __uint_as_float(dp[c][j + 1]), d4.x, d4.y);
            bwd_dsv2(o2, o3, __uint_as_float(s[c][j + 2]), __uint_as_float(s[c][j + 3]), __uint_as_float(dp[c][j + 2]), __uint_as_float(dp[c][j + 3]), d4.z, d4.w);
            pk[c * 16 + j / 2] = bwd_pack<F16>(o0, o1);
            pk[c * 16 + j / 2 + 1] = bwd_pack<F16>(o2, o3);
          }
      } else {
#pragma unroll
        for (int c = 0; c < kBwdC; ++c)
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float ds[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int col = c * 32 + j + e;
              const int pos = i0 + part * kBwdCols + col + shift;
              const bool vis = (p.window_right < 0 || kj <= pos + p.window_right) && (p.window_left < 0 || kj >= pos - p.window_left);
              float dfac = 0.f;
              const float pr = vis ? bwd_prob(p, __uint_as_float(s[c][j + e]), -nl[col], dfac) : 0.f;
              ds[e] = pr * (__uint_as_float(dp[c][j + e]) - dd[col]) * dfac;
            }
            pk[c * 16 + j / 2] = bwd_pack<F16>(ds[0], ds[1]);
          }
      }
      bwd_tmem_st(tB, pk);   // dS^T over the first half of this thread's own dP^T columns
      tmem_st_wait();
      bwd_warp_arrive(&bar->ops_ready[0], lane);
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar->rows_empty[slot]);  // this warp is done with the step's row constants
    }
    if (steps > 0) { mbar_wait(&bar->mma_done, 0, 0xC4); tc_fence_after(); }
    const bool row_ok = kj < sq_.sk;
    const uint32_t tRow = tmem + (uint32_t(q4 * 32) << 16);
    uint16_t* dk_dst = reinterpret_cast<uint16_t*>(p.dk) + sq_.bk * p.dk_batch_stride + int64_t(sq_.k_off + kj) * p.dk_row_stride + hk * p.dk_head_stride;
    uint16_t* dv_dst = reinterpret_cast<uint16_t*>(p.dv) + sq_.bk * p.dv_batch_stride + int64_t(sq_.k_off + kj) * p.dv_row_stride + hk * p.dv_head_stride;
    for (int which = 0; which < 2; ++which) {
      uint16_t* dst = which == 0 ? dk_dst : dv_dst;
      const float sc = which == 0 ? p.scale : 1.0f;
      for (int c = part * kBwdC; c < (part + 1) * kBwdC; ++c) {
        if (c * 32 >= p.d) break;
        uint32_t o[32];
        if (steps > 0) { tmem_ld32(tRow + 256 + which * 128 + c * 32, o); tmem_ld_wait(); }
        else {
#pragma unroll
          for (int j = 0; j < 32; ++j) o[j] = 0u;
        }
        if (row_ok) {
#pragma unroll
          for (int v8 = 0; v8 < 4; ++v8)
            if (c * 32 + v8 * 8 < p.d) {
              uint4 u;
              uint32_t* up = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
              for (int e = 0; e < 4; ++e) up[e] = bwd_pack<F16>(__uint_as_float(o[v8 * 8 + 2 * e]) * sc, __uint_as_float(o[v8 * 8 + 2 * e + 1]) * sc);
              *reinterpret_cast<uint4*>(dst + c * 32 + v8 * 8) = u;
            }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// ----------------------------------------------------------------------------- host
static int launch_bwd(const Fa100BwdParams* p, cudaStream_t stream) {
  FA_CHECK_ARG(p->dout && p->q && p->k && p->v && p->out && p->lse && p->dq && p->dk && p->dv && p->softmax_d, "null tensor");
  FA_CHECK_ARG(p->dtype == FA100_BF16 || p->dtype == FA100_FP16, "FlashAttention backward only supports fp16 and bf16 data type");
  FA_CHECK_ARG(p->b > 0 && p->h > 0 && p->h_k > 0 && p->h % p->h_k == 0, "bad batch/head counts b=%d h=%d h_k=%d", p->b, p->h, p->h_k);
  FA_CHECK_ARG(p->seqlen_q > 0 && p->seqlen_k > 0, "empty sequences");
  FA_CHECK_ARG(p->d > 0 && p->d % 8 == 0, "head dim must be a multiple of 8 (d=%d)", p->d);
  FA_UNSUPPORTED(p->d > 128 || p->dv_dim != p->d, "backward: head dim > 128 or d_v != d is not supported yet (d=%d dv=%d)", p->d, p->dv_dim);
  const bool varlen_q = p->cu_seqlens_q != nullptr, varlen_k = p->cu_seqlens_k != nullptr;
  if (varlen_q) FA_CHECK_ARG(p->total_q > 0, "cu_seqlens_q needs total_q");
  if (varlen_k) FA_CHECK_ARG(p->total_k > 0, "cu_seqlens_k needs total_k");
  // window normalisation of the reference (hopper/flash_api.cpp:1356-1361): a window that covers everything is no window
  int wl = p->window_left, wr = p->window_right;
  if (wl >= p->seqlen_k - 1) wl = -1;
  if (wr >= p->seqlen_q - 1) wr = -1;
  if (p->is_causal) wr = 0;
  for (const int64_t* st : {&p->dq_row_stride, &p->dq_head_stride, &p->dq_batch_stride, &p->dk_row_stride, &p->dk_head_stride, &p->dk_batch_stride,
                            &p->dv_row_stride, &p->dv_head_stride, &p->dv_batch_stride})
    FA_CHECK_ARG(*st % 8 == 0, "dq / dk / dv strides must be multiples of 8 elements");
  FA_CHECK_ARG(reinterpret_cast<uintptr_t>(p->dq) % 16 == 0 && reinterpret_cast<uintptr_t>(p->dk) % 16 == 0 && reinterpret_cast<uintptr_t>(p->dv) % 16 == 0,
               "dq / dk / dv must be 16-byte aligned");

  BwdKernelParams kp;
  std::memset(&kp, 0, sizeof(kp));
  kp.dq = p->dq; kp.dk = p->dk; kp.dv = p->dv; kp.lse = p->lse; kp.dsum = p->softmax_d;
  kp.dq_batch_stride = p->dq_batch_stride; kp.dq_row_stride = p->dq_row_stride; kp.dq_head_stride = p->dq_head_stride;
  kp.dk_batch_stride = p->dk_batch_stride; kp.dk_row_stride = p->dk_row_stride; kp.dk_head_stride = p->dk_head_stride;
  kp.dv_batch_stride = p->dv_batch_stride; kp.dv_row_stride = p->dv_row_stride; kp.dv_head_stride = p->dv_head_stride;
  kp.b = p->b; kp.h = p->h; kp.h_k = p->h_k; kp.seqlen_q = p->seqlen_q; kp.seqlen_k = p->seqlen_k; kp.d = p->d;
  kp.boxes = (p->d + 63) / 64; kp.slices = (p->d + 15) / 16; kp.n16 = (p->d + 15) / 16 * 16;
  kp.f16 = p->dtype == FA100_FP16;
  kp.window_left = wl; kp.window_right = wr;
  kp.scale = p->softmax_scale; kp.scale_log2 = p->softmax_scale * kLog2e;
  if (p->softcap > 0.f) { kp.softcap_pre = p->softmax_scale / p->softcap; kp.softcap_log2 = p->softcap * kLog2e; }
  kp.cu_seqlens_q = p->cu_seqlens_q; kp.cu_seqlens_k = p->cu_seqlens_k; kp.seqused_q = p->seqused_q; kp.seqused_k = p->seqused_k;
  const int64_t rows_q = varlen_q ? p->total_q : p->seqlen_q;
  const int64_t rows_q_rounded = (rows_q + kBwdTile - 1) / kBwdTile * kBwdTile;
  kp.lse_head_stride = rows_q; kp.lse_batch_stride = varlen_q ? 0 : int64_t(p->h) * rows_q;
  kp.dsum_head_stride = rows_q_rounded; kp.dsum_batch_stride = varlen_q ? 0 : int64_t(p->h) * rows_q_rounded;

  CUtensorMap tmQ, tmdO, tmK, tmV;
  const uint32_t box[4] = {64, uint32_t(kBwdTile), 1, 1};
  auto tm4 = [&](CUtensorMap* tm, const void* base, int64_t rows, int heads, int64_t rs, int64_t hs, int64_t bs, bool packed) {
    uint64_t dims[4] = {uint64_t(p->d), uint64_t(rows), uint64_t(heads), uint64_t(packed ? 1 : p->b)};
    uint64_t str[4] = {0, uint64_t(rs) * 2, uint64_t(hs) * 2, uint64_t(bs) * 2};
    return encode_tensor_map(tm, p->dtype, 4, base, dims, str, box);
  };
  int rc;
  const int64_t rows_k = varlen_k ? p->total_k : p->seqlen_k;
  if ((rc = tm4(&tmQ, p->q, rows_q, p->h, p->q_row_stride, p->q_head_stride, p->q_batch_stride, varlen_q))) return rc;
  if ((rc = tm4(&tmdO, p->dout, rows_q, p->h, p->do_row_stride, p->do_head_stride, p->do_batch_stride, varlen_q))) return rc;
  if ((rc = tm4(&tmK, p->k, rows_k, p->h_k, p->k_row_stride, p->k_head_stride, p->k_batch_stride, varlen_k))) return rc;
  if ((rc = tm4(&tmV, p->v, rows_k, p->h_k, p->v_row_stride, p->v_head_stride, p->v_batch_stride, varlen_k))) return rc;

  // the 192 KB of dynamic shared memory are an opt-in per device
  constexpr int kSmemDq = 7 * kBwdTileBytes + 256, kSmemDkdv = 6 * kBwdTileBytes + 2 * kBwdRowSlots * kBwdTile * 4 + 256;
  static std::atomic<uint64_t> attr_set[2] = {{0}, {0}};
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  const bool tracked = dev >= 0 && dev < 128;
  if (!tracked || !((attr_set[dev >> 6].load(std::memory_order_relaxed) >> (dev & 63)) & 1)) {
    FA_CUDA(cudaFuncSetAttribute(fa100_bwd_dq_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemDq));
    FA_CUDA(cudaFuncSetAttribute(fa100_bwd_dq_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemDq));
    FA_CUDA(cudaFuncSetAttribute(fa100_bwd_dkdv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemDkdv));
    FA_CUDA(cudaFuncSetAttribute(fa100_bwd_dkdv_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemDkdv));
    if (tracked) attr_set[dev >> 6].fetch_or(uint64_t(1) << (dev & 63), std::memory_order_relaxed);
  }
  // D: a packed batch is one "sequence" of total_q rows for this kernel (rows no sequence uses are computed and never read)
  const int dot_b = varlen_q ? 1 : p->b, dot_sq = int(rows_q), dot_sqr = int(rows_q_rounded);
  const int64_t rows = int64_t(dot_b) * p->h * dot_sqr;
  const int dot_grid = int(std::min<int64_t>((rows + 7) / 8, 148 * 16));
  const bool dot_vec = reinterpret_cast<uintptr_t>(p->dout) % 16 == 0 && reinterpret_cast<uintptr_t>(p->out) % 16 == 0 && p->do_row_stride % 8 == 0 &&
                       p->do_head_stride % 8 == 0 && p->do_batch_stride % 8 == 0 && p->o_row_stride % 8 == 0 && p->o_head_stride % 8 == 0 &&
                       p->o_batch_stride % 8 == 0;
#define FA_DOT(KERNEL, T)                                                                                                                      \
  FA_CUDA(launch_pdl(KERNEL<T>, dim3(dot_grid), dim3(256), size_t(0), stream, 1, static_cast<const T*>(p->dout), static_cast<const T*>(p->out), \
                     p->softmax_d, p->do_batch_stride, p->do_row_stride, p->do_head_stride, p->o_batch_stride, p->o_row_stride, p->o_head_stride,  \
                     dot_b, p->h, dot_sq, dot_sqr, p->d))
  if (p->dtype == FA100_FP16) { if (dot_vec) FA_DOT(fa100_bwd_dot_vec_kernel, __half); else FA_DOT(fa100_bwd_dot_kernel, __half); }
  else { if (dot_vec) FA_DOT(fa100_bwd_dot_vec_kernel, __nv_bfloat16); else FA_DOT(fa100_bwd_dot_kernel, __nv_bfloat16); }
#undef FA_DOT
  const uint32_t fmt = p->dtype == FA100_BF16 ? kFmtBF16 : kFmtF16;
  const int m_blocks = (p->seqlen_q + kBwdTile - 1) / kBwdTile, n_blocks = (p->seqlen_k + kBwdTile - 1) / kBwdTile;  // of the longest sequence
  const bool f16 = p->dtype == FA100_FP16;
  FA_CUDA(launch_pdl(f16 ? fa100_bwd_dq_kernel<true> : fa100_bwd_dq_kernel<false>, dim3(m_blocks, p->h, p->b), dim3(kBwdThreads), size_t(kSmemDq), stream, 1,
                     kp, tmQ, tmdO, tmK, tmV, fmt));
  FA_CUDA(launch_pdl(f16 ? fa100_bwd_dkdv_kernel<true> : fa100_bwd_dkdv_kernel<false>, dim3(n_blocks, p->h_k, p->b), dim3(kBwdThreads), size_t(kSmemDkdv), stream, 1,
                     kp, tmQ, tmdO, tmK, tmV, fmt));
  FA_CUDA(cudaGetLastError());
  return FA100_OK;
}

}  // namespace fa100

extern "C" int fa100_bwd(const Fa100BwdParams* p, fa100_stream_t stream) {
  if (!p) return fa100::set_error(FA100_ERR_INVALID, "null params");
  return fa100::launch_bwd(p, static_cast<cudaStream_t>(stream));
}
