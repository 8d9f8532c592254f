// fa100_append.cu -- append-KV + rotary pass that precedes the attention kernel.
//
// Reference behaviour (hopper/mainloop_fwd_sm90_tma_gmma_ws.hpp:1447-1713 store_kv_new, :1098-1123 rotary on Q,
// hopper/rotary.h:17-485, hopper/seqlen.h:60-93): K_new (rotated) and V_new are written into the cache in place at rows
// seqused_k[b] .. seqused_k[b] + n_new[b], Q is rotated (a copy; the caller's q is never modified), and attention then
// runs over seqused_k[b] + n_new[b] keys.  Rotary positions: key row i sits at rot_base[b] + i with rot_base =
// seqlens_rotary[b] if given, else seqused_k[b]; query row i sits at rot_base[b] + i when the mask is causal or local and
// at rot_base[b] otherwise (FixedPosition).  Math is fp32, data and cos/sin in the storage dtype (bf16 / fp16 / e4m3), one rounding.
//
// On B200 this is a separate memory-bound pass on the same stream instead of a phase inside the attention kernel: the
// kernel boundary gives the gmem-write -> TMA-read ordering the reference needs a proxy fence for, and the attention
// kernel keeps its TMA-only Q/K/V path.  Bytes moved: |K_new| + |V_new| read and written once, |Q| read and written once
// (only when rotary is on).  One thread moves one 16-byte vector (two for a half-split rotary pair).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>

#include <algorithm>

#include "fa100_host.h"

namespace fa100 {

struct AppendParams {
  const uint8_t* q; uint8_t* q_out;
  uint8_t* k; uint8_t* v;
  const uint8_t* k_new; const uint8_t* v_new;
  const uint8_t* cos; const uint8_t* sin;
  int64_t q_batch_stride, q_row_stride, q_head_stride;
  int64_t qo_batch_stride, qo_row_stride, qo_head_stride;
  int64_t k_batch_stride, k_row_stride, k_head_stride;
  int64_t v_batch_stride, v_row_stride, v_head_stride;
  int64_t knew_batch_stride, knew_row_stride, knew_head_stride;
  int64_t vnew_batch_stride, vnew_row_stride, vnew_head_stride;
  const int32_t* cu_seqlens_q; const int32_t* seqused_q; const int32_t* cu_seqlens_knew;
  const int32_t* seqused_k; const int32_t* kv_batch_idx; const int32_t* seqlens_rotary;
  const int32_t* page_table; int64_t page_table_batch_stride;
  int32_t* seqused_k_out;
  int32_t b, seqlen_q, seqlen_knew, h, h_k, d, dv, page_size, rotary_dim, interleaved, q_moves, elt;
  int32_t cache_rows;  // rows one cache sequence can hold (dense) or page_size * pages per sequence (paged)
};

template <typename T> __device__ __forceinline__ float to_f32(T x);
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }
template <> __device__ __forceinline__ float to_f32<__half>(__half x) { return __half2float(x); }
template <> __device__ __forceinline__ float to_f32<__nv_fp8_e4m3>(__nv_fp8_e4m3 x) { return float(x); }
template <typename T> __device__ __forceinline__ T from_f32(float x);
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }
template <> __device__ __forceinline__ __half from_f32<__half>(float x) { return __float2half_rn(x); }
template <> __device__ __forceinline__ __nv_fp8_e4m3 from_f32<__nv_fp8_e4m3>(float x) { return __nv_fp8_e4m3(x); }  // RN, saturating

// One vector = 8 elements of any storage type: 16 bytes of bf16 / fp16, 8 bytes of e4m3 (the reference's rotary works on
// 16-byte vectors too, hopper/rotary.h:196-267; eight e4m3 values keep rotary_dim % 16 == 0 the only constraint).
constexpr int kVec = 8;
template <typename T> struct VecOf { using Raw = uint4; };
template <> struct VecOf<__nv_fp8_e4m3> { using Raw = uint2; };
template <typename T> union Vec8 { typename VecOf<T>::Raw u; T e[kVec]; __device__ Vec8() {} };

// Moves one head row (`dim` elements at `src`) to `dst`, rotating the first rotary_dim columns at position `pos`.
// Vector j of the row is handled by the calling thread; returns without work for the second vector of a half-split pair.
template <typename T>
__device__ __forceinline__ void move_vec(const AppendParams& a, const uint8_t* src, uint8_t* dst, int j, int pos, bool rotate) {
  using Raw = typename VecOf<T>::Raw;
  const int col = j * kVec;
  const int half = a.rotary_dim / 2;
  if (!rotate || col >= a.rotary_dim) {
    *reinterpret_cast<Raw*>(dst + size_t(col) * sizeof(T)) = *reinterpret_cast<const Raw*>(src + size_t(col) * sizeof(T));
    return;
  }
  const T* cosr = reinterpret_cast<const T*>(a.cos) + size_t(pos) * half;
  const T* sinr = reinterpret_cast<const T*>(a.sin) + size_t(pos) * half;
  if (a.interleaved) {  // pairs (2i, 2i+1): four pairs inside this vector, cos/sin columns col/2 .. col/2+3
    Vec8<T> x; x.u = *reinterpret_cast<const Raw*>(src + size_t(col) * sizeof(T));
#pragma unroll
    for (int i = 0; i < kVec / 2; ++i) {
      const float c = to_f32(cosr[col / 2 + i]), s = to_f32(sinr[col / 2 + i]);
      const float x0 = to_f32(x.e[2 * i]), x1 = to_f32(x.e[2 * i + 1]);
      x.e[2 * i] = from_f32<T>(x0 * c - x1 * s);
      x.e[2 * i + 1] = from_f32<T>(x0 * s + x1 * c);
    }
    *reinterpret_cast<Raw*>(dst + size_t(col) * sizeof(T)) = x.u;
  } else {              // pairs (i, i + rotary_dim/2): this thread owns the vector in the first half and its partner
    if (col >= half) return;
    Vec8<T> lo, hi;
    lo.u = *reinterpret_cast<const Raw*>(src + size_t(col) * sizeof(T));
    hi.u = *reinterpret_cast<const Raw*>(src + size_t(col + half) * sizeof(T));
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      const float c = to_f32(cosr[col + i]), s = to_f32(sinr[col + i]);
      const float x0 = to_f32(lo.e[i]), x1 = to_f32(hi.e[i]);
      lo.e[i] = from_f32<T>(x0 * c - x1 * s);
      hi.e[i] = from_f32<T>(x0 * s + x1 * c);
    }
    *reinterpret_cast<Raw*>(dst + size_t(col) * sizeof(T)) = lo.u;
    *reinterpret_cast<Raw*>(dst + size_t(col + half) * sizeof(T)) = hi.u;
  }
}

// grid (chunks, b); every block walks the vectors of batch element blockIdx.y with a grid-stride loop:
// first the K_new rows, then the V_new rows, then (rotary only) the Q rows.
template <typename T>
__global__ void __launch_bounds__(256) fa100_append_kernel(const AppendParams a) {
  asm volatile("griddepcontrol.wait;\n" ::: "memory");               // PDL (see fa100_ptx.cuh): no global access before this
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  const int bi = blockIdx.y;
  constexpr int vec_elems = kVec;
  const int new0 = a.cu_seqlens_knew ? a.cu_seqlens_knew[bi] : 0;
  const int n_new = a.cu_seqlens_knew ? a.cu_seqlens_knew[bi + 1] - new0 : a.seqlen_knew;
  const int used = a.seqused_k[bi];
  const int rot_base = a.seqlens_rotary ? a.seqlens_rotary[bi] : used;
  const int bk = a.page_table ? 0 : (a.kv_batch_idx ? a.kv_batch_idx[bi] : bi);
  const int kvec = a.d / vec_elems, vvec = a.dv / vec_elems;
  const int64_t n_k = int64_t(n_new) * a.h_k * kvec, n_v = int64_t(n_new) * a.h_k * vvec;
  const int q0 = a.cu_seqlens_q ? a.cu_seqlens_q[bi] : 0;
  const int sq = a.q_moves ? (a.seqused_q ? a.seqused_q[bi] : (a.cu_seqlens_q ? a.cu_seqlens_q[bi + 1] - q0 : a.seqlen_q)) : 0;
  const int64_t n_q = int64_t(sq) * a.h * kvec;
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.seqused_k_out) a.seqused_k_out[bi] = used + n_new;

  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n_k + n_v + n_q; i += int64_t(gridDim.x) * blockDim.x) {
    if (i < n_k + n_v) {
      const bool is_k = i < n_k;
      const int64_t r = is_k ? i : i - n_k;
      const int nvec = is_k ? kvec : vvec;
      const int j = int(r % nvec);
      const int hk = int((r / nvec) % a.h_k);
      const int row = int(r / (int64_t(nvec) * a.h_k));
      const int pos = used + row;  // absolute row in the cache buffer (cache_seqlens counts from the buffer start)
      if (pos >= a.cache_rows) continue;  // the caller promised room (hopper/flash_attn_interface.py:672-680); never write past it
      const uint8_t* src = is_k ? a.k_new + ((a.cu_seqlens_knew ? 0 : bi * a.knew_batch_stride) + int64_t(new0 + row) * a.knew_row_stride + hk * a.knew_head_stride) * a.elt
                                : a.v_new + ((a.cu_seqlens_knew ? 0 : bi * a.vnew_batch_stride) + int64_t(new0 + row) * a.vnew_row_stride + hk * a.vnew_head_stride) * a.elt;
      int64_t cb = bk, cr = pos;
      if (a.page_table) { cb = a.page_table[bi * a.page_table_batch_stride + pos / a.page_size]; cr = pos % a.page_size; }
      uint8_t* dst = is_k ? a.k + (cb * a.k_batch_stride + cr * a.k_row_stride + hk * a.k_head_stride) * a.elt
                          : a.v + (cb * a.v_batch_stride + cr * a.v_row_stride + hk * a.v_head_stride) * a.elt;
      move_vec<T>(a, src, dst, j, rot_base + row, is_k && a.rotary_dim > 0);
    } else {
      const int64_t r = i - n_k - n_v;
      const int j = int(r % kvec);
      const int hq = int((r / kvec) % a.h);
      const int row = int(r / (int64_t(kvec) * a.h));
      const uint8_t* src = a.q + ((a.cu_seqlens_q ? 0 : bi * a.q_batch_stride) + int64_t(q0 + row) * a.q_row_stride + hq * a.q_head_stride) * a.elt;
      uint8_t* dst = a.q_out + ((a.cu_seqlens_q ? 0 : bi * a.qo_batch_stride) + int64_t(q0 + row) * a.qo_row_stride + hq * a.qo_head_stride) * a.elt;
      move_vec<T>(a, src, dst, j, rot_base + (a.q_moves == 2 ? row : 0), true);
    }
  }
}

}  // namespace fa100

extern "C" int fa100_append_kv(const Fa100FwdParams* p, void* q_rotated, int64_t qr_batch_stride, int64_t qr_row_stride,
                               int64_t qr_head_stride, int32_t* seqused_k_out, fa100_stream_t stream_) {
  using namespace fa100;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  FA_CHECK_ARG(p != nullptr, "null params");
  FA_CHECK_ARG(p->k_new && p->v_new, "If k_new is supplied, v_new must also be passed in");
  FA_CHECK_ARG(p->seqused_k != nullptr, "If k_new is supplied, seqlens_k must also be passed in");
  FA_CHECK_ARG(p->cu_seqlens_k == nullptr, "appending to a packed (cu_seqlens_k) cache is not supported");
  FA_CHECK_ARG(p->dtype == FA100_BF16 || p->dtype == FA100_FP16 || p->dtype == FA100_E4M3, "unsupported dtype %d", p->dtype);
  const bool fp8 = p->dtype == FA100_E4M3;
  const int vec = kVec;          // elements per vector
  const int align = fp8 ? 8 : 16;  // bytes per vector
  FA_CHECK_ARG(p->d % vec == 0 && p->dv % vec == 0, "head dims must be multiples of %d", vec);
  const bool rotary = p->rotary_cos != nullptr;
  if (rotary) {
    FA_CHECK_ARG(p->rotary_sin != nullptr, "If rotary cos is provided, rotary sin must also be provided");
    FA_CHECK_ARG(p->rotary_dim > 0 && p->rotary_dim <= p->d, "rotary_dim must be <= headdim");
    FA_CHECK_ARG(p->rotary_dim % 16 == 0, "Only rotary dimensions divisible by 16 are currently supported");
    FA_CHECK_ARG(q_rotated != nullptr && p->q != nullptr, "rotary needs q and a buffer for the rotated q");
  }
  FA_CHECK_ARG(p->knew_row_stride % vec == 0 && p->knew_head_stride % vec == 0 && p->knew_batch_stride % vec == 0 &&
                   p->vnew_row_stride % vec == 0 && p->vnew_head_stride % vec == 0 && p->vnew_batch_stride % vec == 0 &&
                   reinterpret_cast<uintptr_t>(p->k_new) % align == 0 && reinterpret_cast<uintptr_t>(p->v_new) % align == 0,
               "k_new / v_new must be %d-byte aligned with strides that are multiples of %d elements", align, vec);
  FA_CHECK_ARG(p->k_row_stride % vec == 0 && p->k_head_stride % vec == 0 && p->v_row_stride % vec == 0 && p->v_head_stride % vec == 0,
               "the KV cache must have strides that are multiples of %d elements", vec);
  AppendParams a = {};
  a.q = static_cast<const uint8_t*>(p->q); a.q_out = static_cast<uint8_t*>(q_rotated);
  a.k = static_cast<uint8_t*>(const_cast<void*>(p->k)); a.v = static_cast<uint8_t*>(const_cast<void*>(p->v));
  a.k_new = static_cast<const uint8_t*>(p->k_new); a.v_new = static_cast<const uint8_t*>(p->v_new);
  a.cos = static_cast<const uint8_t*>(p->rotary_cos); a.sin = static_cast<const uint8_t*>(p->rotary_sin);
  a.q_batch_stride = p->q_batch_stride; a.q_row_stride = p->q_row_stride; a.q_head_stride = p->q_head_stride;
  a.qo_batch_stride = qr_batch_stride; a.qo_row_stride = qr_row_stride; a.qo_head_stride = qr_head_stride;
  a.k_batch_stride = p->k_batch_stride; a.k_row_stride = p->k_row_stride; a.k_head_stride = p->k_head_stride;
  a.v_batch_stride = p->v_batch_stride; a.v_row_stride = p->v_row_stride; a.v_head_stride = p->v_head_stride;
  a.knew_batch_stride = p->knew_batch_stride; a.knew_row_stride = p->knew_row_stride; a.knew_head_stride = p->knew_head_stride;
  a.vnew_batch_stride = p->vnew_batch_stride; a.vnew_row_stride = p->vnew_row_stride; a.vnew_head_stride = p->vnew_head_stride;
  a.cu_seqlens_q = p->cu_seqlens_q; a.seqused_q = p->seqused_q; a.cu_seqlens_knew = p->cu_seqlens_knew;
  a.seqused_k = p->seqused_k; a.kv_batch_idx = p->kv_batch_idx; a.seqlens_rotary = p->seqlens_rotary;
  a.page_table = p->page_table; a.page_table_batch_stride = p->page_table_batch_stride;
  a.seqused_k_out = seqused_k_out;
  a.b = p->b; a.seqlen_q = p->seqlen_q; a.seqlen_knew = p->seqlen_knew; a.h = p->h; a.h_k = p->h_k; a.d = p->d; a.dv = p->dv;
  a.page_size = p->page_size; a.rotary_dim = rotary ? p->rotary_dim : 0; a.interleaved = p->is_rotary_interleaved;
  // FixedPosition = !(Is_causal || Is_local) AFTER the window normalisation of hopper/flash_api.cpp:796-805 (a window that
  // covers the whole problem is no window: every query row then rotates at the cache length)
  int wl = p->window_left, wr = p->window_right;
  if (wl >= p->seqlen_k - 1) wl = -1;
  if (wr >= p->seqlen_q - 1) wr = -1;
  if (p->is_causal) wr = 0;
  const bool causal = wl < 0 && wr == 0 && p->attention_chunk == 0;
  const bool local = !causal && (wl >= 0 || wr >= 0 || p->attention_chunk > 0);
  a.q_moves = rotary ? ((causal || local) ? 2 : 1) : 0;  // 2: position advances with the row, 1: fixed position
  a.elt = fp8 ? 1 : 2;
  a.cache_rows = p->seqlen_k;
  const int64_t max_new = p->cu_seqlens_knew ? p->total_knew : p->seqlen_knew;
  const int64_t vecs = max_new * p->h_k * ((p->d + p->dv) / vec) + (rotary ? int64_t(p->seqlen_q) * p->h * (p->d / vec) : 0);
  const int chunks = int(std::max<int64_t>(1, std::min<int64_t>((vecs + 255) / 256, 148 * 8 / std::max(1, p->b) + 1)));
  const dim3 grid(chunks, p->b);
  if (fp8) FA_CUDA(launch_pdl(fa100_append_kernel<__nv_fp8_e4m3>, grid, dim3(256), size_t(0), stream, 1, a));
  else if (p->dtype == FA100_FP16) FA_CUDA(launch_pdl(fa100_append_kernel<__half>, grid, dim3(256), size_t(0), stream, 1, a));
  else FA_CUDA(launch_pdl(fa100_append_kernel<__nv_bfloat16>, grid, dim3(256), size_t(0), stream, 1, a));
  FA_CUDA(cudaGetLastError());
  return FA100_OK;
}
