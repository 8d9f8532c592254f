"""Host layer of the forward path: validation, allocation, normalisation and the calls into libfa100.so.

Python counterpart of the reference's ``mha_fwd`` / ``mha_combine`` / ``mha_fwd_get_scheduler_metadata``
(hopper/flash_api.cpp:672-1198, :1569-1670, :521-662).  Registered below as
``torch.ops.flash_attn_3.{fwd,fwd_combine,get_scheduler_metadata,bwd}`` with the reference's frozen schemas
(hopper/flash_api.cpp:1672-1761, pinned by hopper/test_flash_attn.py:1163-1201).  PyTorch is plumbing here:
device memory, the current stream, nothing else -- all arithmetic runs in the CUDA library.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

import fa100_native as N

_DT = {torch.bfloat16: N.BF16, torch.float16: N.FP16, torch.float8_e4m3fn: N.E4M3, torch.float32: N.FP32}


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _chk(cond, msg):
    if not cond:
        raise RuntimeError(msg)


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _i32(t, name, n=None):
    _chk(t.dtype == torch.int32, f"{name} must have dtype torch.int32")
    _chk(t.is_cuda and t.is_contiguous(), f"{name} must be a contiguous CUDA tensor")
    if n is not None:
        _chk(t.numel() == n, f"{name} must have shape ({n},)")


_SM_COUNT = {}
_PLAN_CACHE = {}


def _sm_count(device) -> int:
    idx = torch.cuda.current_device() if device is None else (device.index if isinstance(device, torch.device) and device.index is not None
                                                              else torch.cuda.current_device() if isinstance(device, torch.device) else int(device))
    n = _SM_COUNT.get(idx)
    if n is None:
        n = _SM_COUNT[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return n


def plan(b, seqlen_q, seqlen_k, h, h_k, d, dv, dtype, causal, local, num_splits, pack_gqa, sm_margin, device=None):
    """Sizes-only policy (num_splits, pack_gqa, #work items); the role of get_num_splits/get_pack_gqa.  A pure function of
    its arguments, so results are cached (a decode loop asks the same question every step; measured host cost of a call:
    profiles/r02_host_latency_*.json)."""
    num_sm = max(1, _sm_count(device) - sm_margin)
    key = (b, seqlen_q, seqlen_k, h, h_k, d, dv, dtype, bool(causal), bool(local), int(num_splits), pack_gqa, num_sm)
    p = _PLAN_CACHE.get(key)
    if p is None:
        s = N.Shape(b=b, seqlen_q=seqlen_q, seqlen_k=seqlen_k, h=h, h_k=h_k, d=d, dv=dv, dtype=_DT[dtype],
                    is_causal=int(causal), is_local=int(local), num_splits=int(num_splits),
                    pack_gqa=-1 if pack_gqa is None else int(bool(pack_gqa)), num_sm=num_sm)
        p = N.Plan()
        N.check(N.lib().fa100_plan(C.byref(s), C.byref(p)))
        if len(_PLAN_CACHE) > 4096:
            _PLAN_CACHE.clear()
        _PLAN_CACHE[key] = p
    return p


class _on_device:
    """torch.cuda.device(dev) only when dev is not already current (the context manager costs ~5 us of host time per call)."""

    def __init__(self, dev):
        self.ctx = None if dev.index is None or dev.index == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            return self.ctx.__exit__(*exc)
        return False


def mha_fwd(q, k, v, k_new=None, v_new=None, q_v=None, out=None, cu_seqlens_q=None, cu_seqlens_k=None,
            cu_seqlens_k_new=None, seqused_q=None, seqused_k=None, max_seqlen_q=None, max_seqlen_k=None,
            page_table=None, kv_batch_idx=None, leftpad_k=None, rotary_cos=None, rotary_sin=None,
            seqlens_rotary=None, q_descale=None, k_descale=None, v_descale=None, softmax_scale=None,
            is_causal=False, window_size_left=-1, window_size_right=-1, attention_chunk=0, softcap=0.0,
            is_rotary_interleaved=False, scheduler_metadata=None, num_splits=0, pack_gqa=None, sm_margin=0, _partial=None, _merge=0):
    """torch.ops.flash_attn_3.fwd.  `_merge` (with `_partial`): 1 = merge this call's result into the partial already in the
    buffers, in place; 2 = merge and write `out` / the returned lse in the output dtype (Fa100FwdParams.merge_partial).  `_partial=(o_accum, lse_accum)` (not part of the schema; used by ring_attention) makes
    the kernel write one fp32 partial -- o_accum (1, b, h, sq, dv), lse_accum (1, b, h, sq) -- instead of out / lse."""
    _chk(q.is_cuda and k.is_cuda and v.is_cuda, "q, k, v must be CUDA tensors (there is no CPU path)")
    qt = q.dtype
    _chk(qt in (torch.bfloat16, torch.float16, torch.float8_e4m3fn), "FlashAttention only supports fp16, bf16, and fp8_e4m3 data type")
    _chk(k.dtype == qt and v.dtype == qt, "query, key and value must have the same dtype")
    _chk(q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1, "Input tensor must have contiguous last dimension")
    paged = page_table is not None
    varlen_q = cu_seqlens_q is not None
    varlen_k = cu_seqlens_k is not None
    _chk(not (paged and varlen_k), "If cu_seqlens_k is passed in, then page table is not supported")
    if varlen_q:
        _i32(cu_seqlens_q, "cu_seqlens_q")
        _chk(max_seqlen_q is not None, "max_seqlen_q must be provided if cu_seqlens_q is provided")
        _chk(q.dim() == 3, "q must be (total_q, nheads, headdim) when cu_seqlens_q is given")
    if varlen_k:
        _i32(cu_seqlens_k, "cu_seqlens_k")
        _chk(max_seqlen_k is not None, "max_seqlen_k must be provided if cu_seqlens_k is provided")
        _chk(k.dim() == 3, "k must be (total_k, nheads_k, headdim) when cu_seqlens_k is given")
    b = (cu_seqlens_q.numel() - 1) if varlen_q else q.shape[0]
    seqlen_q = max_seqlen_q if varlen_q else q.shape[1]
    total_q = q.shape[0] if varlen_q else 0
    h, d = q.shape[-2], q.shape[-1]
    h_k, dv = k.shape[-2], v.shape[-1]
    if paged:
        _i32(page_table, "page_table")
        _chk(page_table.dim() == 2 and page_table.shape[0] == b, "page_table must have shape (batch_size, max_num_pages_per_seq)")
        num_pages, page_size = k.shape[0], k.shape[1]
        seqlen_k = page_table.shape[1] * page_size
        b_k = num_pages
        _chk(tuple(v.shape) == (num_pages, page_size, h_k, dv), "v has the wrong paged shape")
    else:
        num_pages = page_size = 0
        seqlen_k = max_seqlen_k if varlen_k else k.shape[1]
        b_k = (cu_seqlens_k.numel() - 1) if varlen_k else k.shape[0]
        if not varlen_k:
            _chk(tuple(v.shape) == (b_k, seqlen_k, h_k, dv), "v has the wrong shape")
    total_k = k.shape[0] if varlen_k else 0
    _chk(k.shape[-1] == d, "k must have the same head dim as q")
    _chk(h % h_k == 0, "Number of heads in key/value must divide number of heads in query")
    if kv_batch_idx is None and not varlen_k and not paged:
        _chk(b_k == b, "batch_size must be equal to batch_size_k")
    for t, nm in ((seqused_q, "seqused_q"), (seqused_k, "seqused_k"), (leftpad_k, "leftpad_k"), (kv_batch_idx, "kv_batch_idx")):
        if t is not None:
            _i32(t, nm, b)
    if leftpad_k is not None:
        _chk(not paged, "leftpad_k is not supported with paged KV")
    if q_v is not None:  # hopper/flash_api.cpp:1028-1050
        _chk(d <= 64, "q_v is only supported for head_size <= 64")
        _chk(dv >= 256, "q_v is only supported for hdim_v >= 256.")
        _chk(qt in (torch.float16, torch.bfloat16), "q_v is only supported for fp16 and bf16 data type")
        _chk(q_v.dtype == qt, "q_v must have the same dtype as query")
        _chk(q_v.is_cuda and q_v.stride(-1) == 1, "q_v tensor must have contiguous last dimension")
        _chk(tuple(q_v.shape) == ((total_q, h, dv) if varlen_q else (b, seqlen_q, h, dv)), "q_v has the wrong shape")
        if dv > 256:
            raise NotImplementedError("fa100: q_v with hdim_v > 256 is not supported")
    align = 16 if qt == torch.float8_e4m3fn else 8
    _chk(d % align == 0 and dv % align == 0, f"head dims must be multiples of {align} for {qt}")
    if rotary_cos is not None:
        _chk(k_new is not None, "If rotary cos/sin are provided, new key / value to be appended to KV cache must also be provided")
    if k_new is not None:  # append-KV (+ rotary) pass, hopper/flash_api.cpp:929-975, 1052-1082
        q, seqused_k = _append_kv(q, k, v, k_new, v_new, cu_seqlens_q, cu_seqlens_k_new, seqused_q, seqused_k, page_table,
                                  kv_batch_idx, rotary_cos, rotary_sin, seqlens_rotary, is_rotary_interleaved, is_causal,
                                  window_size_left, window_size_right, attention_chunk, b, seqlen_q, seqlen_k, h, h_k, d, dv,
                                  varlen_q, varlen_k, paged, page_size)
    if softmax_scale is None:
        softmax_scale = d ** -0.5
    if is_causal:
        window_size_right = 0
    is_local = (window_size_left >= 0 or window_size_right >= 0 or attention_chunk > 0) and not (
        window_size_left < 0 and window_size_right == 0 and attention_chunk == 0)

    out_dtype = torch.bfloat16 if qt == torch.float8_e4m3fn else qt
    out_shape = (total_q, h, dv) if varlen_q else (b, seqlen_q, h, dv)
    if _partial is not None and _merge != 2:  # the kernel writes only the fp32 partial
        out = torch.empty((0,) * len(out_shape), dtype=out_dtype, device=q.device)
    elif out is None:
        out = torch.empty(out_shape, dtype=out_dtype, device=q.device)
    else:
        _chk(out.dtype == out_dtype and tuple(out.shape) == out_shape and out.stride(-1) == 1 and out.device == q.device,
             "out has the wrong dtype / shape / layout")
    lse = torch.empty(((h, total_q) if varlen_q else (b, h, seqlen_q)) if (_partial is None or _merge == 2) else (0,), dtype=torch.float32, device=q.device)
    empty = torch.empty(0, dtype=torch.float32, device=q.device)
    if _partial is None and (seqlen_k == 0 or out.numel() == 0):  # hopper/flash_api.cpp:1190-1194
        out.zero_()
        lse.fill_(float("inf"))
        return out, lse, empty, empty
    if scheduler_metadata is not None:
        _chk(scheduler_metadata.dtype == torch.int32 and scheduler_metadata.is_cuda and scheduler_metadata.is_contiguous(),
             "scheduler_metadata must be a contiguous int32 CUDA tensor")

    if _partial is not None:
        num_splits = 1
    pl = plan(b, seqlen_q, seqlen_k, h, h_k, d, dv, qt, is_causal, is_local, num_splits, pack_gqa, sm_margin, q.device)
    ragged = b > 1 and any(t is not None for t in (cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k, leftpad_k)) and not _NO_SCHED
    if ragged and num_splits == 0 and pl.dual_stream and pl.num_splits < DYN_SPLIT_BOUND and scheduler_metadata is None and _partial is None:
        # ragged decode: the schedule kernel picks a split count per batch element (long sequences are cut, short ones are
        # not; hopper/flash_prepare_scheduler.cu:96-120); the plan only sets the bound the partial buffers are sized for
        pl = plan(b, seqlen_q, seqlen_k, h, h_k, d, dv, qt, is_causal, is_local, DYN_SPLIT_BOUND, pack_gqa, sm_margin, q.device)
    splits = pl.num_splits
    out_accum = lse_accum = None
    if _partial is not None:
        out_accum, lse_accum = _partial
        _chk(not varlen_q and out_accum.dtype == torch.float32 and out_accum.is_contiguous() and lse_accum.is_contiguous()
             and tuple(out_accum.shape) == (1, b, h, seqlen_q, dv) and tuple(lse_accum.shape) == (1, b, h, seqlen_q),
             "_partial buffers must be contiguous fp32 (1, b, h, sq, dv) / (1, b, h, sq)")
    elif splits > 1:
        if varlen_q:
            out_accum = torch.empty((splits, h, total_q, dv), dtype=torch.float32, device=q.device)
            lse_accum = torch.empty((splits, h, total_q), dtype=torch.float32, device=q.device)
        else:
            out_accum = torch.empty((splits, b, h, seqlen_q, dv), dtype=torch.float32, device=q.device)
            lse_accum = torch.empty((splits, b, h, seqlen_q), dtype=torch.float32, device=q.device)

    p = N.FwdParams()
    p.q, p.k, p.v, p.o, p.lse = _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(lse)
    if varlen_q:
        p.q_batch_stride, p.q_row_stride, p.q_head_stride = 0, q.stride(0), q.stride(1)
        p.o_batch_stride, p.o_row_stride, p.o_head_stride = 0, out.stride(0), out.stride(1)
    else:
        p.q_batch_stride, p.q_row_stride, p.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
        p.o_batch_stride, p.o_row_stride, p.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    if varlen_k:
        p.k_batch_stride, p.k_row_stride, p.k_head_stride = 0, k.stride(0), k.stride(1)
        p.v_batch_stride, p.v_row_stride, p.v_head_stride = 0, v.stride(0), v.stride(1)
    else:
        p.k_batch_stride, p.k_row_stride, p.k_head_stride = k.stride(0), k.stride(1), k.stride(2)
        p.v_batch_stride, p.v_row_stride, p.v_head_stride = v.stride(0), v.stride(1), v.stride(2)
    if q_v is not None:
        p.qv = _ptr(q_v)
        p.qv_batch_stride, p.qv_row_stride, p.qv_head_stride = ((0, q_v.stride(0), q_v.stride(1)) if varlen_q else
                                                                (q_v.stride(0), q_v.stride(1), q_v.stride(2)))
    p.o_accum, p.lse_accum = _ptr(out_accum), _ptr(lse_accum)
    p.b, p.b_k, p.seqlen_q, p.seqlen_k, p.total_q, p.total_k = b, b_k, seqlen_q, seqlen_k, total_q, total_k
    p.h, p.h_k, p.d, p.dv, p.dtype = h, h_k, d, dv, _DT[qt]
    p.cu_seqlens_q, p.cu_seqlens_k = _ptr(cu_seqlens_q), _ptr(cu_seqlens_k)
    p.seqused_q, p.seqused_k, p.leftpad_k, p.kv_batch_idx = _ptr(seqused_q), _ptr(seqused_k), _ptr(leftpad_k), _ptr(kv_batch_idx)
    if paged:
        p.page_table, p.page_table_batch_stride = _ptr(page_table), page_table.stride(0)
        p.page_size, p.num_pages = page_size, num_pages
    for nm, t in (("q", q_descale), ("k", k_descale), ("v", v_descale)):
        if t is not None:
            _chk(qt == torch.float8_e4m3fn or True, "")
            _chk(t.dtype == torch.float32 and t.is_cuda and tuple(t.shape) == (b, h_k), f"{nm}_descale must be fp32 (batch, nheads_k)")
            setattr(p, f"{nm}_descale", _ptr(t))
            setattr(p, f"{nm}_descale_batch_stride", t.stride(0))
            setattr(p, f"{nm}_descale_head_stride", t.stride(1))
    p.softmax_scale, p.softcap = float(softmax_scale), float(softcap)
    p.is_causal, p.window_left, p.window_right, p.attention_chunk = int(is_causal), window_size_left, window_size_right, attention_chunk
    p.num_splits, p.pack_gqa, p.num_sm = splits, pl.pack_gqa, pl.num_sm
    p.variant = _variant()
    p.merge_partial = int(_merge) if _partial is not None else 0
    with _on_device(q.device):
        counter = None
        use_sched = (scheduler_metadata is not None and scheduler_metadata.numel() >= META_HEADER + _sched_words(b)) or ragged
        if _DYNAMIC and pl.num_work > pl.num_sm and (use_sched or ((is_causal or window_size_left >= 0 or window_size_right >= 0) and seqlen_q > 256)):
            # dynamic tile scheduling (the reference's tile_count_semaphore, zeroed per call) when the items are unequal (masks, ragged
            # batches) and there are more of them than CTAs: a fresh counter per launch, so calls on different streams or replays of
            # different graphs never share one.  Zeroed BEFORE the schedule kernel so that kernel and the forward stay PDL-chained.
            counter = torch.zeros(1, dtype=torch.int32, device=q.device)
            p.tile_counter = _ptr(counter)
        if scheduler_metadata is not None and scheduler_metadata.numel() >= META_HEADER + _sched_words(b):
            # the schedule get_scheduler_metadata built ahead of time (CUDA-graphed decode): used as is
            _chk(int(scheduler_metadata.numel()) >= META_HEADER, "scheduler_metadata is too short")
            p.scheduler_metadata = _ptr(scheduler_metadata)
        elif ragged:
            # the reference launches prepare_varlen_num_blocks in front of every ragged forward (hopper/flash_api.cpp:1166)
            meta = torch.empty(META_HEADER + _sched_words(b), dtype=torch.int32, device=q.device)
            N.check(N.lib().fa100_prepare_scheduler(C.byref(p), _ptr(meta), meta.numel(), _stream()))
            p.scheduler_metadata = _ptr(meta)
            p.lse_accum_prefilled = int(lse_accum is not None and splits > 1 and p.merge_partial == 0)  # filled by that launch
        N.check(N.lib().fa100_fwd(C.byref(p), _stream()))
        del counter  # allocated and used on the current stream: the caching allocator reuses it only for later work of that stream
        if _partial is None and splits > 1:
            if varlen_q:  # (splits, h, total_q, dv) viewed as (splits, b=1, s=total_q, h, dv)
                op = out_accum.permute(0, 2, 1, 3).unsqueeze(1)
                lp = lse_accum.permute(0, 2, 1).unsqueeze(1)
                _combine_into(op, lp, out.unsqueeze(0), lse.t().unsqueeze(0))
            else:
                _combine_into(out_accum.permute(0, 1, 3, 2, 4), lse_accum.permute(0, 1, 3, 2), out, lse.permute(0, 2, 1), seqused=seqused_q)
    return out, lse, (out_accum if out_accum is not None else empty), (lse_accum if lse_accum is not None else empty)


META_HEADER = 8  # host-written words in front of the device-built schedule
DYN_SPLIT_BOUND = 8  # partials a ragged decode call is sized for when the caller leaves num_splits to the library
_DYNAMIC = os.environ.get("FA100_STATIC_SCHEDULE") != "1"  # A/B switch: FA100_STATIC_SCHEDULE=1 = no tile counter (static round-robin)
_NO_SCHED = bool(os.environ.get("FA100_NO_SCHED"))  # A/B switch: static schedule for ragged batches too


def _variant() -> int:
    """Fa100FwdParams.variant: FA100_PAIR=1 forces the 2-CTA (cta_group::2) kernel wherever it is eligible, FA100_PAIR=0
    forbids it; unset, the library's default applies (one CTA per work item: the pair kernel measures within 1 % of it)."""
    v = {"1": 1, "0": 2}.get(os.environ.get("FA100_PAIR", ""), 0)
    if os.environ.get("FA100_NO_TMA_STORE"):   # A/B switches live in the Python layer; the C library reads no environment
        v |= 16
    if os.environ.get("FA100_FOUR_ROLE") == "1":  # the four-role prefill mainloop (csrc/fa100_fwd2.cuh)
        v |= 32
    if os.environ.get("FA100_SPLIT_ROWS") == "1":  # two softmax threads per query row (csrc/fa100_fwd_split.cuh)
        v |= 64
    return v


def _sched_words(b: int) -> int:
    return int(N.lib().fa100_scheduler_metadata_words(b)) - META_HEADER


def _append_kv(q, k, v, k_new, v_new, cu_seqlens_q, cu_seqlens_k_new, seqused_q, seqused_k, page_table, kv_batch_idx,
               rotary_cos, rotary_sin, seqlens_rotary, interleaved, is_causal, wl, wr, chunk, b, seqlen_q, seqlen_k, h, h_k, d, dv,
               varlen_q, varlen_k, paged, page_size):
    """Write k_new (rotated) / v_new into the cache in place, rotate a copy of q, return (q', cache_seqlens + n_new)."""
    qt = q.dtype
    _chk(v_new is not None, "If k_new is supplied, v_new must also be passed in")
    _chk(seqused_k is not None, "If k_new is supplied, seqlens_k must also be passed in")
    _chk(seqlen_q <= seqlen_k, "If k_new is supplied, it must have seqlen <= the seqlen of the KV cache")
    _chk(not varlen_k, "appending to a packed (cu_seqlens_k) cache is not supported")
    _chk(k_new.dtype == qt and v_new.dtype == qt, "k_new must have the same dtype as query")
    _chk(k_new.is_cuda and v_new.is_cuda, "k_new, v_new must be CUDA tensors")
    _chk(k_new.stride(-1) == 1 and v_new.stride(-1) == 1, "k_new tensor must have contiguous last dimension")
    varlen_new = cu_seqlens_k_new is not None
    pa = N.FwdParams()
    if varlen_new:
        _i32(cu_seqlens_k_new, "cu_seqlens_k_new", b + 1)
        total_new = k_new.shape[0]
        _chk(tuple(k_new.shape) == (total_new, h_k, d) and tuple(v_new.shape) == (total_new, h_k, dv), "k_new / v_new have the wrong shape")
        pa.knew_batch_stride, pa.knew_row_stride, pa.knew_head_stride = 0, k_new.stride(0), k_new.stride(1)
        pa.vnew_batch_stride, pa.vnew_row_stride, pa.vnew_head_stride = 0, v_new.stride(0), v_new.stride(1)
        pa.seqlen_knew, pa.total_knew, pa.cu_seqlens_knew = 0, total_new, _ptr(cu_seqlens_k_new)
    else:
        s_new = k_new.shape[1]
        _chk(tuple(k_new.shape) == (b, s_new, h_k, d) and tuple(v_new.shape) == (b, s_new, h_k, dv), "k_new / v_new have the wrong shape")
        pa.knew_batch_stride, pa.knew_row_stride, pa.knew_head_stride = k_new.stride(0), k_new.stride(1), k_new.stride(2)
        pa.vnew_batch_stride, pa.vnew_row_stride, pa.vnew_head_stride = v_new.stride(0), v_new.stride(1), v_new.stride(2)
        pa.seqlen_knew, pa.total_knew = s_new, b * s_new
    pa.k_new, pa.v_new = _ptr(k_new), _ptr(v_new)
    pa.q, pa.k, pa.v = _ptr(q), _ptr(k), _ptr(v)
    if varlen_q:
        pa.q_batch_stride, pa.q_row_stride, pa.q_head_stride = 0, q.stride(0), q.stride(1)
    else:
        pa.q_batch_stride, pa.q_row_stride, pa.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
    pa.k_batch_stride, pa.k_row_stride, pa.k_head_stride = k.stride(0), k.stride(1), k.stride(2)
    pa.v_batch_stride, pa.v_row_stride, pa.v_head_stride = v.stride(0), v.stride(1), v.stride(2)
    pa.b, pa.seqlen_q, pa.seqlen_k, pa.h, pa.h_k, pa.d, pa.dv, pa.dtype = b, seqlen_q, seqlen_k, h, h_k, d, dv, _DT[qt]
    pa.cu_seqlens_q, pa.seqused_q, pa.seqused_k, pa.kv_batch_idx = _ptr(cu_seqlens_q), _ptr(seqused_q), _ptr(seqused_k), _ptr(kv_batch_idx)
    if paged:
        pa.page_table, pa.page_table_batch_stride, pa.page_size = _ptr(page_table), page_table.stride(0), page_size
    pa.is_causal, pa.window_left, pa.window_right, pa.attention_chunk = int(is_causal), wl, wr, chunk
    q_rot = None
    if rotary_cos is not None:
        _chk(rotary_sin is not None, "If rotary cos is provided, rotary sin must also be provided")
        _chk(rotary_cos.is_cuda and rotary_cos.is_contiguous() and rotary_sin.is_cuda and rotary_sin.is_contiguous(),
             "rotary_cos / rotary_sin must be contiguous CUDA tensors")
        rotary_dim = rotary_cos.shape[1] * 2
        _chk(rotary_dim <= d, "rotary_dim must be <= headdim")
        _chk(rotary_dim % 16 == 0, "Only rotary dimensions divisible by 16 are currently supported")
        _chk(tuple(rotary_sin.shape) == tuple(rotary_cos.shape), "rotary_sin must have the shape of rotary_cos")
        _chk(rotary_cos.dtype == qt and rotary_sin.dtype == qt, "rotary_cos must have the same dtype as query")
        if seqlens_rotary is not None:
            _i32(seqlens_rotary, "seqlens_rotary", b)
        pa.rotary_cos, pa.rotary_sin, pa.seqlens_rotary = _ptr(rotary_cos), _ptr(rotary_sin), _ptr(seqlens_rotary)
        pa.rotary_dim, pa.is_rotary_interleaved = rotary_dim, int(bool(interleaved))
        q_rot = torch.empty(q.shape, dtype=qt, device=q.device)
    new_lens = torch.empty(b, dtype=torch.int32, device=q.device)
    qs = (0, q_rot.stride(0), q_rot.stride(1)) if (q_rot is not None and varlen_q) else (
        (q_rot.stride(0), q_rot.stride(1), q_rot.stride(2)) if q_rot is not None else (0, 0, 0))
    with _on_device(q.device):
        N.check(N.lib().fa100_append_kv(C.byref(pa), _ptr(q_rot), qs[0], qs[1], qs[2], _ptr(new_lens), _stream()))
    return (q_rot if q_rot is not None else q), new_lens


def _combine_into(op, lp, out, lse, seqused=None):
    """op (splits,b,s,h,d) fp32, lp (splits,b,s,h) fp32, out (b,s,h,d), lse (b,s,h) -- any strides, d stride 1."""
    splits, b, s, h, d = op.shape
    c = N.CombineParams()
    c.o_partial, c.lse_partial, c.o, c.lse = _ptr(op), _ptr(lp), _ptr(out), _ptr(lse)
    c.op_split_stride, c.op_batch_stride, c.op_row_stride, c.op_head_stride = op.stride(0), op.stride(1), op.stride(2), op.stride(3)
    c.lp_split_stride, c.lp_batch_stride, c.lp_row_stride, c.lp_head_stride = lp.stride(0), lp.stride(1), lp.stride(2), lp.stride(3)
    c.o_batch_stride, c.o_row_stride, c.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    c.lse_batch_stride, c.lse_row_stride, c.lse_head_stride = lse.stride(0), lse.stride(1), lse.stride(2)
    c.num_splits, c.b, c.s, c.h, c.d, c.out_dtype = splits, b, s, h, d, _DT[out.dtype]
    if seqused is not None:  # rows past a batch element's used length keep what the caller's buffer held
        c.seqused = _ptr(seqused)
    N.check(N.lib().fa100_combine(C.byref(c), _stream()))


def mha_combine(out_partial, lse_partial, out=None, out_dtype=None):
    """hopper/flash_api.cpp:1569-1670: out_partial (splits,b,s,h,d) fp32, lse_partial (splits,b,s,h) fp32."""
    _chk(out_partial.is_cuda and lse_partial.is_cuda, "inputs must be CUDA tensors")
    _chk(out_partial.dtype == torch.float32 and lse_partial.dtype == torch.float32, "FlashAttention combine only supports fp32 partials")
    _chk(out_partial.dim() == 5 and lse_partial.dim() == 4, "out_partial must be 5-D and lse_partial 4-D")
    _chk(out_partial.stride(-1) == 1, "Input tensor must have contiguous last dimension")
    splits, b, s, h, d = out_partial.shape
    _chk(tuple(lse_partial.shape) == (splits, b, s, h), "lse_partial has the wrong shape")
    _chk(1 <= splits <= 256, "FlashAttention combine only supports num_splits at most 256")
    if out_dtype is None:
        out_dtype = out.dtype if out is not None else torch.float32
    _chk(out_dtype in (torch.float32, torch.float16, torch.bfloat16), "out_dtype must be fp32, fp16 or bf16")
    d4 = (d + 3) // 4 * 4
    op = out_partial
    if d4 != d or any(st % 4 for st in op.stride()[:-1]) or op.data_ptr() % 16:
        op = torch.nn.functional.pad(out_partial, (0, d4 - d)).contiguous()
    if out is None:
        out_p = torch.empty((b, s, h, d4), dtype=out_dtype, device=op.device)
    else:
        _chk(tuple(out.shape) == (b, s, h, d) and out.dtype == out_dtype and out.stride(-1) == 1, "out has the wrong shape/dtype")
        out_p = out if d4 == d and all(st % 4 == 0 for st in out.stride()[:-1]) else torch.empty((b, s, h, d4), dtype=out_dtype, device=op.device)
    lse = torch.empty((b, h, s), dtype=torch.float32, device=op.device).transpose(1, 2)  # (b, s, h) view, s contiguous
    with _on_device(op.device):
        _combine_into(op, lse_partial, out_p, lse)
    if out is None:
        out = out_p[..., :d] if d4 != d else out_p
    elif out_p is not out:
        out.copy_(out_p[..., :d])
    return out, lse


def mha_fwd_get_scheduler_metadata(batch_size, max_seqlen_q, max_seqlen_k, num_heads, num_heads_k, headdim, headdim_v,
                                   qkv_dtype, seqused_k, cu_seqlens_q=None, cu_seqlens_k=None, cu_seqlens_k_new=None,
                                   seqused_q=None, leftpad_k=None, page_size=None, max_seqlen_k_new=0, is_causal=False,
                                   window_size_left=-1, window_size_right=-1, attention_chunk=0, has_softcap=False,
                                   num_splits=0, pack_gqa=None, sm_margin=0):
    """Ahead-of-time plan for CUDA-graphed decode (hopper/flash_api.cpp:521-662).  The fa100 schedule is a pure
    function of sizes, so the tensor only records it: [magic, num_splits, pack_gqa, num_sm, num_work, 0, 0, 0]."""
    if is_causal:
        window_size_right = 0
    is_local = (window_size_left >= 0 or window_size_right >= 0 or attention_chunk > 0) and not (
        window_size_left < 0 and window_size_right == 0 and attention_chunk == 0)
    dev = seqused_k.device
    pl = plan(batch_size, max_seqlen_q, max_seqlen_k + max_seqlen_k_new, num_heads, num_heads_k, headdim, headdim_v, qkv_dtype,
              is_causal, is_local, num_splits, pack_gqa, sm_margin, dev)
    header = torch.tensor([0xFA100, pl.num_splits, pl.pack_gqa, pl.num_sm, pl.num_work, 0, 0, 0], dtype=torch.int32, device=dev)
    if batch_size <= 1 or cu_seqlens_k_new is not None or max_seqlen_k_new > 0:
        # one sequence has nothing to order; with appended keys the lengths change before the forward reads them
        return header
    # ragged batch: the device-built list of work items (role of prepare_varlen_num_blocks, hopper/flash_api.cpp:640-660)
    meta = torch.empty(META_HEADER + _sched_words(batch_size), dtype=torch.int32, device=dev)
    meta[:META_HEADER] = header
    p = N.FwdParams()
    p.b, p.b_k, p.seqlen_q, p.seqlen_k = batch_size, batch_size, max_seqlen_q, max_seqlen_k
    p.h, p.h_k, p.d, p.dv, p.dtype = num_heads, num_heads_k, headdim, headdim_v, _DT[qkv_dtype]
    p.cu_seqlens_q, p.cu_seqlens_k = _ptr(cu_seqlens_q), _ptr(cu_seqlens_k)
    p.seqused_q, p.seqused_k, p.leftpad_k = _ptr(seqused_q), _ptr(seqused_k), _ptr(leftpad_k)
    p.total_q = 1 if cu_seqlens_q is not None else 0  # sizes of the packed tensors play no part in the schedule
    p.is_causal, p.window_left, p.window_right, p.attention_chunk = int(is_causal), window_size_left, window_size_right, attention_chunk
    p.num_splits, p.pack_gqa, p.num_sm = pl.num_splits, pl.pack_gqa, pl.num_sm
    p.variant = _variant()
    with _on_device(dev):
        N.check(N.lib().fa100_prepare_scheduler(C.byref(p), _ptr(meta), meta.numel(), _stream()))
    return meta


def mha_bwd(dout, q, k, v, out, softmax_lse, dq=None, dk=None, dv=None, cu_seqlens_q=None, cu_seqlens_k=None, seqused_q=None,
            seqused_k=None, max_seqlen_q=None, max_seqlen_k=None, softmax_scale=None, is_causal=False, window_size_left=-1,
            window_size_right=-1, softcap=0.0, deterministic=False, sm_margin=0):
    """torch.ops.flash_attn_3.bwd (hopper/flash_api.cpp:1267-1567): bf16 / fp16, d == dv <= 128, MHA / GQA / MQA, fixed-length or
    variable-length (cu_seqlens / seqused) batches, causal, local windows, softcap; larger head dims raise NotImplementedError.
    Returns (dq, dk, dv, softmax_d, softmax_lse_log2, dq_accum, dk_accum, dv_accum) like the reference; the last four are
    empty (no fp32 accumulators: dQ and dK/dV each come from one deterministic kernel that folds log2(e) itself)."""
    _chk(q.is_cuda and k.is_cuda and v.is_cuda and dout.is_cuda and out.is_cuda and softmax_lse.is_cuda, "all tensors must be CUDA tensors")
    qt = q.dtype
    _chk(qt in (torch.float16, torch.bfloat16), "FlashAttention only support fp16 and bf16 data type")
    _chk(k.dtype == qt and v.dtype == qt and out.dtype == qt and dout.dtype == qt, "query, key, value, out and dout must have the same dtype")
    for t, nm in ((q, "q"), (k, "k"), (v, "v"), (out, "out"), (dout, "dout")):
        _chk(t.stride(-1) == 1, f"{nm} tensor must have contiguous last dimension")
    varlen_q, varlen_k = cu_seqlens_q is not None, cu_seqlens_k is not None
    if varlen_q:
        _i32(cu_seqlens_q, "cu_seqlens_q")
        _chk(max_seqlen_q is not None, "max_seqlen_q must be provided if cu_seqlens_q is provided")
        _chk(q.dim() == 3, "q must be (total_q, nheads, headdim) when cu_seqlens_q is given")
    if varlen_k:
        _i32(cu_seqlens_k, "cu_seqlens_k")
        _chk(max_seqlen_k is not None, "max_seqlen_k must be provided if cu_seqlens_k is provided")
        _chk(k.dim() == 3, "k must be (total_k, nheads_k, headdim) when cu_seqlens_k is given")
    b = (cu_seqlens_q.numel() - 1) if varlen_q else q.shape[0]
    sq = max_seqlen_q if varlen_q else q.shape[1]
    sk = max_seqlen_k if varlen_k else k.shape[1]
    total_q, total_k = (q.shape[0] if varlen_q else 0), (k.shape[0] if varlen_k else 0)
    h, d = q.shape[-2], q.shape[-1]
    h_k, dv_dim = v.shape[-2], v.shape[-1]
    _chk(tuple(k.shape) == ((total_k, h_k, d) if varlen_k else (b, sk, h_k, d)), "k has the wrong shape")
    _chk(tuple(v.shape) == ((total_k, h_k, dv_dim) if varlen_k else (b, sk, h_k, dv_dim)), "v has the wrong shape")
    oshape = (total_q, h, dv_dim) if varlen_q else (b, sq, h, dv_dim)
    _chk(tuple(out.shape) == oshape and tuple(dout.shape) == oshape, "out / dout have the wrong shape")
    _chk(h % h_k == 0, "Number of heads in key/value must divide number of heads in query")
    _chk(softmax_lse.dtype == torch.float32 and tuple(softmax_lse.shape) == ((h, total_q) if varlen_q else (b, h, sq)),
         "softmax_lse must be fp32 (batch, nheads, seqlen_q) or (nheads, total_q)")
    _chk(d % 8 == 0 and dv_dim % 8 == 0, "head_size should be a multiple of 8")
    for t, nm in ((seqused_q, "seqused_q"), (seqused_k, "seqused_k")):
        if t is not None:
            _i32(t, nm, b)
    if softmax_scale is None:
        softmax_scale = d ** -0.5
    mk = lambda ref, t: (torch.empty_like(ref) if t is None else t)  # noqa: E731
    dq, dk, dv = mk(q, dq), mk(k, dk), mk(v, dv)
    for t, ref, nm in ((dq, q, "dq"), (dk, k, "dk"), (dv, v, "dv")):
        _chk(t.dtype == qt and tuple(t.shape) == tuple(ref.shape) and t.stride(-1) == 1 and t.is_cuda, f"{nm} must have the dtype, shape and layout of its input")
    if seqused_q is not None or seqused_k is not None:
        # rows past seqused are never visited: the reference zero-fills them too (dq / dk / dv are defined for every slot)
        dq.zero_(), dk.zero_(), dv.zero_()
    lse = softmax_lse.contiguous()
    rows_q = total_q if varlen_q else sq
    rows_q_rounded = (rows_q + 127) // 128 * 128
    softmax_d = torch.empty((h, rows_q_rounded) if varlen_q else (b, h, rows_q_rounded), dtype=torch.float32, device=q.device)
    p = N.BwdParams()
    p.dout, p.q, p.k, p.v, p.out, p.lse = _ptr(dout), _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(lse)
    p.dq, p.dk, p.dv, p.softmax_d = _ptr(dq), _ptr(dk), _ptr(dv), _ptr(softmax_d)
    for nm, t, packed in (("q", q, varlen_q), ("k", k, varlen_k), ("v", v, varlen_k), ("o", out, varlen_q), ("do", dout, varlen_q),
                          ("dq", dq, varlen_q), ("dk", dk, varlen_k), ("dv", dv, varlen_k)):
        st = (0, t.stride(0), t.stride(1)) if packed else (t.stride(0), t.stride(1), t.stride(2))
        setattr(p, f"{nm}_batch_stride", st[0])
        setattr(p, f"{nm}_row_stride", st[1])
        setattr(p, f"{nm}_head_stride", st[2])
    p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d, p.dv_dim, p.dtype = b, sq, sk, h, h_k, d, dv_dim, _DT[qt]
    p.cu_seqlens_q, p.cu_seqlens_k, p.seqused_q, p.seqused_k = _ptr(cu_seqlens_q), _ptr(cu_seqlens_k), _ptr(seqused_q), _ptr(seqused_k)
    p.total_q, p.total_k = total_q, total_k
    p.is_causal, p.window_left, p.window_right = int(is_causal), window_size_left, window_size_right
    p.softmax_scale, p.softcap, p.deterministic = float(softmax_scale), float(softcap), int(deterministic)
    p.num_sm = max(1, _sm_count(q.device) - sm_margin)
    with _on_device(q.device):
        N.check(N.lib().fa100_bwd(C.byref(p), _stream()))
    empty = torch.empty(0, dtype=torch.float32, device=q.device)
    return dq, dk, dv, softmax_d, empty, empty, empty, empty


# ----------------------------------------------------------------------------- shape-only (Meta) kernels for tracing
def mha_fwd_meta(q, k, v, k_new=None, v_new=None, q_v=None, out=None, cu_seqlens_q=None, cu_seqlens_k=None,
                 cu_seqlens_k_new=None, seqused_q=None, seqused_k=None, max_seqlen_q=None, max_seqlen_k=None,
                 page_table=None, kv_batch_idx=None, leftpad_k=None, rotary_cos=None, rotary_sin=None,
                 seqlens_rotary=None, q_descale=None, k_descale=None, v_descale=None, softmax_scale=None,
                 is_causal=False, window_size_left=-1, window_size_right=-1, attention_chunk=0, softcap=0.0,
                 is_rotary_interleaved=False, scheduler_metadata=None, num_splits=0, pack_gqa=None, sm_margin=0):
    """Output shapes / dtypes of flash_attn_3::fwd for FakeTensor / torch.compile tracing (the
    reference registers no Meta kernel, hopper/flash_api.cpp:1763-1768).  Partials are only materialised for an explicit
    num_splits > 1; with num_splits = 0 the real kernel may return non-empty accumulators."""
    varlen_q = cu_seqlens_q is not None
    h, dv = q.shape[-2], v.shape[-1]
    out_dtype = torch.bfloat16 if q.dtype == torch.float8_e4m3fn else q.dtype
    if varlen_q:
        total_q = q.shape[0]
        out_shape, lse_shape = (total_q, h, dv), (h, total_q)
        acc_shape, lacc_shape = (num_splits, h, total_q, dv), (num_splits, h, total_q)
    else:
        b, sq = q.shape[0], q.shape[1]
        out_shape, lse_shape = (b, sq, h, dv), (b, h, sq)
        acc_shape, lacc_shape = (num_splits, b, h, sq, dv), (num_splits, b, h, sq)
    o = out if out is not None else q.new_empty(out_shape, dtype=out_dtype)
    lse = q.new_empty(lse_shape, dtype=torch.float32)
    if num_splits > 1:
        return o, lse, q.new_empty(acc_shape, dtype=torch.float32), q.new_empty(lacc_shape, dtype=torch.float32)
    return o, lse, q.new_empty((0,), dtype=torch.float32), q.new_empty((0,), dtype=torch.float32)


def mha_combine_meta(out_partial, lse_partial, out=None, out_dtype=None):
    splits, b, s, h, d = out_partial.shape
    o = out if out is not None else out_partial.new_empty((b, s, h, d), dtype=out_dtype or out_partial.dtype)
    return o, out_partial.new_empty((b, h, s), dtype=torch.float32).transpose(1, 2)


# ----------------------------------------------------------------------------- torch.library registration
_FWD_SCHEMA = (
    "(Tensor q, Tensor k, Tensor v, Tensor(k_new!)? k_new = None, Tensor(v_new!)? v_new = None, Tensor? q_v = None, "
    "Tensor(out!)? out = None, Tensor? cu_seqlens_q = None, Tensor? cu_seqlens_k = None, Tensor? cu_seqlens_k_new = None, "
    "Tensor? seqused_q = None, Tensor? seqused_k = None, int? max_seqlen_q = None, int? max_seqlen_k = None, "
    "Tensor? page_table = None, Tensor? kv_batch_idx = None, Tensor? leftpad_k = None, Tensor? rotary_cos = None, "
    "Tensor? rotary_sin = None, Tensor? seqlens_rotary = None, Tensor? q_descale = None, Tensor? k_descale = None, "
    "Tensor? v_descale = None, float? softmax_scale = None, bool is_causal = False, int window_size_left = -1, "
    "int window_size_right = -1, int attention_chunk = 0, float softcap = 0.0, bool is_rotary_interleaved = False, "
    "Tensor? scheduler_metadata = None, int num_splits = 0, bool? pack_gqa = None, int sm_margin = 0) "
    "-> (Tensor(out!), Tensor, Tensor, Tensor)")
_BWD_SCHEMA = (
    "(Tensor dout, Tensor q, Tensor k, Tensor v, Tensor out, Tensor softmax_lse, Tensor(dq!)? dq = None, "
    "Tensor(dk!)? dk = None, Tensor(dv!)? dv = None, Tensor? cu_seqlens_q = None, Tensor? cu_seqlens_k = None, "
    "Tensor? seqused_q = None, Tensor? seqused_k = None, int? max_seqlen_q = None, int? max_seqlen_k = None, "
    "float? softmax_scale = None, bool is_causal = False, int window_size_left = -1, int window_size_right = -1, "
    "float softcap = 0.0, bool deterministic = False, int sm_margin = 0) "
    "-> (Tensor(dq!), Tensor(dk!), Tensor(dv!), Tensor, Tensor, Tensor, Tensor, Tensor)")
_COMBINE_SCHEMA = "(Tensor out_partial, Tensor lse_partial, Tensor(out!)? out = None, ScalarType? out_dtype = None) -> (Tensor(out!), Tensor)"
_META_SCHEMA = (
    "(int batch_size, int max_seqlen_q, int max_seqlen_k, int num_heads, int num_heads_k, int headdim, int headdim_v, "
    "ScalarType qkv_dtype, Tensor seqused_k, Tensor? cu_seqlens_q = None, Tensor? cu_seqlens_k = None, "
    "Tensor? cu_seqlens_k_new = None, Tensor? seqused_q = None, Tensor? leftpad_k = None, int? page_size = None, "
    "int max_seqlen_k_new = 0, bool is_causal = False, int window_size_left = -1, int window_size_right = -1, "
    "int attention_chunk = 0, bool has_softcap = False, int num_splits = 0, bool? pack_gqa = None, int sm_margin = 0) -> Tensor")

_lib_def = None


def register_ops():
    """Define flash_attn_3::* once per process: CUDA kernels like the reference (no CPU kernel), plus shape-only Meta
    kernels so that FakeTensor / torch.compile can trace through the ops."""
    global _lib_def
    if _lib_def is not None:
        return
    if hasattr(torch.ops, "flash_attn_3") and hasattr(torch.ops.flash_attn_3, "fwd"):
        raise RuntimeError("torch.ops.flash_attn_3 is already defined by another extension in this process")
    _lib_def = torch.library.Library("flash_attn_3", "DEF")
    _lib_def.define("fwd" + _FWD_SCHEMA)
    _lib_def.define("bwd" + _BWD_SCHEMA)
    _lib_def.define("fwd_combine" + _COMBINE_SCHEMA)
    _lib_def.define("get_scheduler_metadata" + _META_SCHEMA)
    _lib_def.impl("fwd", mha_fwd, "CUDA")
    _lib_def.impl("bwd", mha_bwd, "CUDA")
    _lib_def.impl("fwd_combine", mha_combine, "CUDA")
    _lib_def.impl("get_scheduler_metadata", mha_fwd_get_scheduler_metadata, "CUDA")
    _lib_def.impl("fwd", mha_fwd_meta, "Meta")
    _lib_def.impl("fwd_combine", mha_combine_meta, "Meta")
