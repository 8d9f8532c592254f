"""Drop-in ``flash_attn_interface`` for the forward path, backed by the B200-native libfa100.so.

Same public names, argument order, defaults and return values as the reference's
hopper/flash_attn_interface.py (``flash_attn_func`` :507, ``flash_attn_varlen_func`` :588,
``flash_attn_qkvpacked_func`` :445, ``flash_attn_combine`` :636, ``flash_attn_with_kvcache`` :640,
``get_scheduler_metadata`` :803).  Forward only: ``backward`` raises NotImplementedError.
"""
from __future__ import annotations

from typing import Optional, Union

import torch

import fa100_ops

fa100_ops.register_ops()
flash_attn_3_cuda = torch.ops.flash_attn_3


def maybe_contiguous(x):
    return x.contiguous() if x is not None and x.stride(-1) != 1 else x


def _flash_attn_forward(q, k, v, k_new, v_new, qv, out, cu_seqlens_q, cu_seqlens_k, cu_seqlens_k_new, seqused_q,
                        seqused_k, max_seqlen_q, max_seqlen_k, page_table, kv_batch_idx, leftpad_k, rotary_cos,
                        rotary_sin, seqlens_rotary, q_descale, k_descale, v_descale, softmax_scale, causal,
                        window_size=(-1, -1), attention_chunk=0, softcap=0.0, rotary_interleaved=True,
                        scheduler_metadata=None, num_splits=1, pack_gqa=None, sm_margin=0):
    q, k, k_new, v_new = [maybe_contiguous(x) for x in (q, k, k_new, v_new)]
    v = v.contiguous() if v.stride(-1) != 1 and v.stride(-3) != 1 else v
    v = maybe_contiguous(v)  # fa100 has no V-column-major path
    cu_seqlens_q, cu_seqlens_k, cu_seqlens_k_new, seqused_q, seqused_k, page_table, kv_batch_idx, leftpad_k, \
        rotary_cos, rotary_sin, seqlens_rotary = [
            maybe_contiguous(x) for x in (cu_seqlens_q, cu_seqlens_k, cu_seqlens_k_new, seqused_q, seqused_k,
                                          page_table, kv_batch_idx, leftpad_k, rotary_cos, rotary_sin, seqlens_rotary)]
    out, softmax_lse, *rest = flash_attn_3_cuda.fwd(
        q, k, v, k_new, v_new, qv, out, cu_seqlens_q, cu_seqlens_k, cu_seqlens_k_new, seqused_q, seqused_k,
        max_seqlen_q, max_seqlen_k, page_table, kv_batch_idx, leftpad_k, rotary_cos, rotary_sin, seqlens_rotary,
        q_descale, k_descale, v_descale, softmax_scale, causal, window_size[0], window_size[1], attention_chunk,
        softcap, rotary_interleaved, scheduler_metadata, num_splits, pack_gqa, sm_margin)
    return out, softmax_lse, *rest


def _flash_attn_backward(dout, q, k, v, out, softmax_lse, cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k, max_seqlen_q,
                         max_seqlen_k, dq, dk, dv, softmax_scale, causal, window_size=(-1, -1), softcap=0.0, deterministic=False,
                         sm_margin=0):
    """hopper/flash_attn_interface.py:119-168."""
    dout, q, k, v, out = [maybe_contiguous(x) for x in (dout, q, k, v, out)]
    dq, dk, dv, softmax_d, *rest = torch.ops.flash_attn_3.bwd(
        dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k, max_seqlen_q, max_seqlen_k,
        softmax_scale, causal, window_size[0], window_size[1], softcap, deterministic, sm_margin)
    return dq, dk, dv, softmax_d


class FlashAttnFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, softmax_scale, causal, qv=None, q_descale=None, k_descale=None, v_descale=None,
                window_size=(-1, -1), attention_chunk=0, softcap=0.0, num_splits=1, pack_gqa=None,
                deterministic=False, sm_margin=0, return_softmax=False):
        if softmax_scale is None:
            softmax_scale = (q.shape[-1] + (qv.shape[-1] if qv is not None else 0)) ** (-0.5)
        out, softmax_lse, *rest = _flash_attn_forward(
            q, k, v, None, None, qv, None, None, None, None, None, None, None, None, None, None, None, None, None,
            None, q_descale, k_descale, v_descale, softmax_scale, causal=causal, window_size=window_size,
            attention_chunk=attention_chunk, softcap=softcap, num_splits=num_splits, pack_gqa=pack_gqa,
            sm_margin=sm_margin)
        ctx.save_for_backward(q, k, v, out, softmax_lse)
        ctx.softmax_scale, ctx.causal, ctx.window_size, ctx.attention_chunk = softmax_scale, causal, window_size, attention_chunk
        ctx.softcap, ctx.deterministic, ctx.sm_margin, ctx.has_qv = softcap, deterministic, sm_margin, qv is not None
        ctx.mark_non_differentiable(softmax_lse)
        return (out, softmax_lse) if return_softmax else out

    @staticmethod
    def backward(ctx, dout, *args):
        """hopper/flash_attn_interface.py:236-264."""
        q, k, v, out, softmax_lse = ctx.saved_tensors
        assert ctx.attention_chunk == 0, "FA3 backward does not support attention_chunk"
        if ctx.has_qv or q.dtype == torch.float8_e4m3fn:
            raise NotImplementedError("fa100: the backward does not support qv or FP8 (neither does the reference's)")
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _flash_attn_backward(dout, q, k, v, out, softmax_lse, None, None, None, None, None, None, dq, dk, dv, ctx.softmax_scale,
                             ctx.causal, ctx.window_size, ctx.softcap, ctx.deterministic, ctx.sm_margin)
        return (dq, dk, dv) + (None,) * 14


class FlashAttnVarlenFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k, max_seqlen_q, max_seqlen_k,
                softmax_scale, causal, qv=None, q_descale=None, k_descale=None, v_descale=None,
                window_size=(-1, -1), attention_chunk=0, softcap=0.0, num_splits=1, pack_gqa=None,
                deterministic=False, sm_margin=0, return_softmax=False):
        if softmax_scale is None:
            softmax_scale = (q.shape[-1] + (qv.shape[-1] if qv is not None else 0)) ** (-0.5)
        out, softmax_lse, *rest = _flash_attn_forward(
            q, k, v, None, None, qv, None, cu_seqlens_q, cu_seqlens_k, None, seqused_q, seqused_k, max_seqlen_q,
            max_seqlen_k, None, None, None, None, None, None, q_descale, k_descale, v_descale, softmax_scale,
            causal=causal, window_size=window_size, attention_chunk=attention_chunk, softcap=softcap,
            num_splits=num_splits, pack_gqa=pack_gqa, sm_margin=sm_margin)
        ctx.save_for_backward(q, k, v, out, softmax_lse, cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k)
        ctx.max_seqlen_q, ctx.max_seqlen_k = max_seqlen_q, max_seqlen_k
        ctx.softmax_scale, ctx.causal, ctx.window_size, ctx.attention_chunk = softmax_scale, causal, window_size, attention_chunk
        ctx.softcap, ctx.deterministic, ctx.sm_margin, ctx.has_qv = softcap, deterministic, sm_margin, qv is not None
        ctx.mark_non_differentiable(softmax_lse)
        return (out, softmax_lse) if return_softmax else out

    @staticmethod
    def backward(ctx, dout, *args):
        """hopper/flash_attn_interface.py:335-367."""
        q, k, v, out, softmax_lse, cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k = ctx.saved_tensors
        assert ctx.attention_chunk == 0, "FA3 backward does not support attention_chunk"
        if ctx.has_qv or q.dtype == torch.float8_e4m3fn:
            raise NotImplementedError("fa100: the backward does not support qv or FP8 (neither does the reference's)")
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _flash_attn_backward(dout, q, k, v, out, softmax_lse, cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k, ctx.max_seqlen_q,
                             ctx.max_seqlen_k, dq, dk, dv, ctx.softmax_scale, ctx.causal, ctx.window_size, ctx.softcap, ctx.deterministic,
                             ctx.sm_margin)
        return (dq, dk, dv) + (None,) * 20


def flash_attn_qkvpacked_func(qkv, softmax_scale=None, causal=False, q_descale=None, k_descale=None, v_descale=None,
                              window_size=(-1, -1), attention_chunk=0, softcap=0.0, deterministic=False,
                              num_heads_q=None, sm_margin=0, return_attn_probs=False):
    """qkv: (batch, seqlen, 3, nheads, headdim), or (batch, seqlen, nheads_q + 2 * nheads_k, headdim) with num_heads_q."""
    if qkv.dim() == 5:
        q, k, v = qkv.unbind(dim=-3)
    else:
        assert qkv.dim() == 4 and num_heads_q is not None
        num_heads_k = (qkv.shape[2] - num_heads_q) // 2
        q, k, v = qkv.split([num_heads_q, num_heads_k, num_heads_k], dim=-2)
    return FlashAttnFunc.apply(q, k, v, softmax_scale, causal, None, q_descale, k_descale, v_descale, window_size,
                               attention_chunk, softcap, 1, None, deterministic, sm_margin, return_attn_probs)


def flash_attn_func(q, k, v, softmax_scale=None, causal=False, qv=None, q_descale=None, k_descale=None, v_descale=None,
                    window_size=(-1, -1), attention_chunk=0, softcap=0.0, num_splits=1, pack_gqa=None,
                    deterministic=False, sm_margin=0, return_attn_probs=False):
    """q (b, sq, h, d), k (b, sk, h_k, d), v (b, sk, h_k, dv) -> out (b, sq, h, dv) [, softmax_lse (b, h, sq)].
    MQA/GQA: q head i reads kv head i // (h / h_k).  causal is aligned to the bottom-right corner;
    window_size=(l, r) keeps keys in [i + sk - sq - l, i + sk - sq + r]; rows with no visible key give 0."""
    return FlashAttnFunc.apply(q, k, v, softmax_scale, causal, qv, q_descale, k_descale, v_descale, window_size,
                               attention_chunk, softcap, num_splits, pack_gqa, deterministic, sm_margin,
                               return_attn_probs)


def flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, seqused_q=None,
                           seqused_k=None, softmax_scale=None, causal=False, qv=None, q_descale=None, k_descale=None,
                           v_descale=None, window_size=(-1, -1), attention_chunk=0, softcap=0.0, num_splits=1,
                           pack_gqa=None, deterministic=False, sm_margin=0, return_attn_probs=False):
    """Packed variable-length batch: q (total_q, h, d), k/v (total_k, h_k, d) with int32 cu_seqlens (b + 1)."""
    return FlashAttnVarlenFunc.apply(q, k, v, cu_seqlens_q, cu_seqlens_k, seqused_q, seqused_k, max_seqlen_q,
                                     max_seqlen_k, softmax_scale, causal, qv, q_descale, k_descale, v_descale,
                                     window_size, attention_chunk, softcap, num_splits, pack_gqa, deterministic,
                                     sm_margin, return_attn_probs)


def flash_attn_combine(out_partial, lse_partial, out=None, out_dtype=None):
    return flash_attn_3_cuda.fwd_combine(out_partial, lse_partial, out, out_dtype)


def flash_attn_with_kvcache(q, k_cache, v_cache, k=None, v=None, qv=None, rotary_cos=None, rotary_sin=None,
                            cache_seqlens: Optional[Union[int, torch.Tensor]] = None,
                            cache_batch_idx: Optional[torch.Tensor] = None, cache_leftpad: Optional[torch.Tensor] = None,
                            page_table: Optional[torch.Tensor] = None, cu_seqlens_q: Optional[torch.Tensor] = None,
                            cu_seqlens_k_new: Optional[torch.Tensor] = None, max_seqlen_q: Optional[int] = None,
                            rotary_seqlens: Optional[torch.Tensor] = None, q_descale: Optional[torch.Tensor] = None,
                            k_descale: Optional[torch.Tensor] = None, v_descale: Optional[torch.Tensor] = None,
                            softmax_scale=None, causal=False, window_size=(-1, -1), attention_chunk=0, softcap=0.0,
                            rotary_interleaved=True, scheduler_metadata=None, num_splits=0, pack_gqa=None, sm_margin=0,
                            return_softmax_lse=False):
    """Decode / prefill against a (paged) KV cache; see the reference docstring (hopper/flash_attn_interface.py:672-757).
    k_cache (b_cache, s_cache, h_k, d) or (num_pages, page_size, h_k, d) with page_table (b, max_pages)."""
    assert k_cache.stride(-1) == 1, "k_cache must have contiguous last dimension"
    assert v_cache.stride(-1) == 1, "v_cache must have contiguous last dimension"
    if softmax_scale is None:
        softmax_scale = (q.shape[-1] + (qv.shape[-1] if qv is not None else 0)) ** (-0.5)
    if cache_seqlens is not None and isinstance(cache_seqlens, int):
        cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=k_cache.device)
    out, softmax_lse, *rest = _flash_attn_forward(
        q, k_cache, v_cache, k, v, qv, None, cu_seqlens_q, None, cu_seqlens_k_new, None, cache_seqlens, max_seqlen_q,
        None, page_table, cache_batch_idx, cache_leftpad, rotary_cos, rotary_sin, rotary_seqlens, q_descale, k_descale,
        v_descale, softmax_scale, causal=causal, window_size=window_size, attention_chunk=attention_chunk,
        softcap=softcap, rotary_interleaved=rotary_interleaved, scheduler_metadata=scheduler_metadata,
        num_splits=num_splits, pack_gqa=pack_gqa, sm_margin=sm_margin)
    return (out, softmax_lse, *rest) if return_softmax_lse else out


def get_scheduler_metadata(batch_size, max_seqlen_q, max_seqlen_k, num_heads_q, num_heads_kv, headdim,
                           cache_seqlens: torch.Tensor, qkv_dtype=torch.bfloat16, headdim_v=None,
                           cu_seqlens_q: Optional[torch.Tensor] = None, cu_seqlens_k_new: Optional[torch.Tensor] = None,
                           cache_leftpad: Optional[torch.Tensor] = None, page_size: Optional[int] = None,
                           max_seqlen_k_new=0, causal=False, window_size=(-1, -1), attention_chunk=0,
                           has_softcap=False, num_splits=0, pack_gqa=None, sm_margin=0):
    cache_seqlens = maybe_contiguous(cache_seqlens)
    if headdim_v is None:
        headdim_v = headdim
    return flash_attn_3_cuda.get_scheduler_metadata(
        batch_size, max_seqlen_q, max_seqlen_k, num_heads_q, num_heads_kv, headdim, headdim_v, qkv_dtype,
        cache_seqlens, cu_seqlens_q, None, cu_seqlens_k_new, None, cache_leftpad, page_size, max_seqlen_k_new, causal,
        window_size[0], window_size[1], attention_chunk, has_softcap, num_splits, pack_gqa, sm_margin)
