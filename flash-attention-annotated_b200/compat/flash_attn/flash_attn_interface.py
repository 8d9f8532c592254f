"""The FA2 public API (/root/reference/flash_attn/flash_attn_interface.py:1003-1590) on top of the fa100 forward.

Same function names, argument order, defaults and return conventions as the reference's FA2 module, so that callers
written against `flash_attn.flash_attn_interface` (the model zoo's `flash_attn/modules/mha.py`, serving stacks that pass
`block_table`) run on the B200 kernel unchanged.  What maps where:

  block_table                 -> page_table (+ seqused_k from cu_seqlens_k in the varlen form)
  cache_seqlens as an int     -> a (batch,) int32 tensor
  head dims not multiple of 8 -> zero-padded to the next multiple of 8 and sliced back (flash_attn_interface.py:465-470)
  return_attn_probs=True      -> (out, softmax_lse, None): S_dmask only exists with dropout (a testing aid there too)
  dropout_p > 0, alibi_slopes -> NotImplementedError (the forward kernel has neither; inference callers pass 0.0 / None)
  deterministic               -> accepted (the forward is deterministic; the flag concerns the backward)

Forward only: like the rest of this package the autograd backward raises."""
from __future__ import annotations

import os
import sys
from typing import Optional, Union

import torch

_PKG = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _PKG not in sys.path:
    sys.path.insert(0, _PKG)
import flash_attn_interface as _fa3  # noqa: E402  (the FA3-style interface of this package; raises if libfa100.so is missing)


def _unsupported(dropout_p, alibi_slopes):
    if dropout_p != 0.0:
        raise NotImplementedError("fa100: dropout_p > 0 is not supported (forward-only inference kernel)")
    if alibi_slopes is not None:
        raise NotImplementedError("fa100: alibi_slopes is not supported by the FA3-style forward")


def _pad8(*ts):
    """FA2 pads head dims that are not a multiple of 8 with zeros (flash_attn_interface.py:465-470)."""
    d = ts[0].shape[-1]
    if d % 8 == 0:
        return ts, d
    return tuple(torch.nn.functional.pad(t, [0, 8 - d % 8]) for t in ts), d


def _ret(out, lse, return_attn_probs):
    return (out, lse, None) if return_attn_probs else out


def flash_attn_func(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False, window_size=(-1, -1), softcap=0.0,
                    alibi_slopes=None, deterministic=False, return_attn_probs=False):
    _unsupported(dropout_p, alibi_slopes)
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    (q, k, v), d = _pad8(q, k, v)
    out, lse = _fa3.flash_attn_func(q, k, v, softmax_scale=softmax_scale, causal=causal, window_size=tuple(window_size),
                                    softcap=softcap, deterministic=deterministic, return_attn_probs=True)
    return _ret(out[..., :d], lse, return_attn_probs)


def flash_attn_kvpacked_func(q, kv, dropout_p=0.0, softmax_scale=None, causal=False, window_size=(-1, -1), softcap=0.0,
                             alibi_slopes=None, deterministic=False, return_attn_probs=False):
    """kv: (batch, seqlen_k, 2, nheads_k, headdim)."""
    return flash_attn_func(q, kv[:, :, 0], kv[:, :, 1], dropout_p, softmax_scale, causal, window_size, softcap, alibi_slopes,
                           deterministic, return_attn_probs)


def flash_attn_qkvpacked_func(qkv, dropout_p=0.0, softmax_scale=None, causal=False, window_size=(-1, -1), softcap=0.0,
                              alibi_slopes=None, deterministic=False, return_attn_probs=False):
    """qkv: (batch, seqlen, 3, nheads, headdim)."""
    return flash_attn_func(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], dropout_p, softmax_scale, causal, window_size, softcap,
                           alibi_slopes, deterministic, return_attn_probs)


def flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p=0.0, softmax_scale=None,
                           causal=False, window_size=(-1, -1), softcap=0.0, alibi_slopes=None, deterministic=False,
                           return_attn_probs=False, block_table=None):
    _unsupported(dropout_p, alibi_slopes)
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    (q, k, v), d = _pad8(q, k, v)
    if block_table is None:
        out, lse = _fa3.flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                                               softmax_scale=softmax_scale, causal=causal, window_size=tuple(window_size), softcap=softcap,
                                               deterministic=deterministic, return_attn_probs=True)
    else:
        # paged K/V (num_blocks, page_size, nheads_k, headdim) addressed by block_table (batch, max_blocks); the key lengths
        # come from cu_seqlens_k (flash_attn_interface.py:1480-1483).  FA3 form: packed q + page_table + per-sequence lengths
        seqused_k = (cu_seqlens_k[1:] - cu_seqlens_k[:-1]).to(torch.int32).contiguous()
        out, lse, *_ = _fa3.flash_attn_with_kvcache(q, k, v, cache_seqlens=seqused_k, page_table=block_table.to(torch.int32),
                                                    cu_seqlens_q=cu_seqlens_q, max_seqlen_q=max_seqlen_q, softmax_scale=softmax_scale,
                                                    causal=causal, window_size=tuple(window_size), softcap=softcap, return_softmax_lse=True)
    return _ret(out[..., :d], lse, return_attn_probs)


def flash_attn_varlen_kvpacked_func(q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p=0.0, softmax_scale=None,
                                    causal=False, window_size=(-1, -1), softcap=0.0, alibi_slopes=None, deterministic=False,
                                    return_attn_probs=False):
    """kv: (total_k, 2, nheads_k, headdim)."""
    return flash_attn_varlen_func(q, kv[:, 0], kv[:, 1], cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                                  softmax_scale, causal, window_size, softcap, alibi_slopes, deterministic, return_attn_probs)


def flash_attn_varlen_qkvpacked_func(qkv, cu_seqlens, max_seqlen, dropout_p=0.0, softmax_scale=None, causal=False,
                                     window_size=(-1, -1), softcap=0.0, alibi_slopes=None, deterministic=False, return_attn_probs=False):
    """qkv: (total, 3, nheads, headdim)."""
    return flash_attn_varlen_func(qkv[:, 0], qkv[:, 1], qkv[:, 2], cu_seqlens, cu_seqlens, max_seqlen, max_seqlen, dropout_p,
                                  softmax_scale, causal, window_size, softcap, alibi_slopes, deterministic, return_attn_probs)


def flash_attn_with_kvcache(q, k_cache, v_cache, k=None, v=None, rotary_cos=None, rotary_sin=None,
                            cache_seqlens: Optional[Union[int, torch.Tensor]] = None, cache_batch_idx: Optional[torch.Tensor] = None,
                            cache_leftpad: Optional[torch.Tensor] = None, block_table: Optional[torch.Tensor] = None, softmax_scale=None,
                            causal=False, window_size=(-1, -1), softcap=0.0, rotary_interleaved=True, alibi_slopes=None, num_splits=0,
                            return_softmax_lse=False):
    """FA2's incremental-decoding entry point (flash_attn_interface.py:1474-1590): k / v are appended to the cache in place at
    cache_seqlens (which is NOT updated), rotary is applied to q and k with cos / sin of shape (seqlen_ro, rotary_dim / 2)."""
    _unsupported(0.0, alibi_slopes)
    assert k_cache.stride(-1) == 1 and v_cache.stride(-1) == 1, "k_cache and v_cache must have contiguous last dimension"
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** (-0.5)
    if cache_seqlens is not None and isinstance(cache_seqlens, int):
        cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=k_cache.device)
    out, lse, *_ = _fa3.flash_attn_with_kvcache(
        q, k_cache, v_cache, k=k, v=v, rotary_cos=rotary_cos, rotary_sin=rotary_sin, cache_seqlens=cache_seqlens,
        cache_batch_idx=cache_batch_idx, cache_leftpad=cache_leftpad, page_table=None if block_table is None else block_table.to(torch.int32),
        softmax_scale=softmax_scale, causal=causal, window_size=tuple(window_size), softcap=softcap, rotary_interleaved=rotary_interleaved,
        num_splits=num_splits, return_softmax_lse=True)
    return (out, lse) if return_softmax_lse else out
