"""CPU oracle for the FlashAttention forward path.  TEST INFRASTRUCTURE ONLY.

This file is a restatement, in plain PyTorch on the CPU, of the reference's
own correctness oracle and of the forward-path semantics collected in
SURVEY.md section 7.0.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the
product path (``flash_attn_interface`` -> C-ABI -> CUDA) never does.

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the reference's
``attention_ref`` (hopper/test_util.py:226-348) and ``attention_combine_ref``
(hopper/test_flash_attn.py:1105-1114) inside the build container and stores
seeded input/output pairs under ``tests/golden/``; ``tests/test_oracle.py``
checks this restatement against every stored pair.

What each function follows:

* ``attention_oracle``    hopper/test_util.py:226-348 (``attention_ref``) plus
  the mask builders at :157-223; additionally returns the log-sum-exp the
  reference tests only carry as a comment (hopper/test_flash_attn.py:183-190).
* ``combine_oracle``      hopper/test_flash_attn.py:1105-1114.
* ``rotary_oracle``       flash_attn/layers/rotary.py semantics as used by
  hopper/test_flash_attn.py:790-830 (interleaved / NeoX-half, per-batch offset).
* ``paged_kvcache``       hopper/test_flash_attn.py:992-1015.
* ``fwd_tolerance``       hopper/test_flash_attn.py:193-223 and :986-989.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

NEG_INF = float("-inf")


# --------------------------------------------------------------------------- masks
def visibility_mask(
    seqlen_q: int,
    seqlen_k: int,
    *,
    causal: bool = False,
    window_size: Tuple[int, int] = (-1, -1),
    attention_chunk: int = 0,
    sink_token_length: int = 0,
    q_lens: Optional[torch.Tensor] = None,  # (b,) valid query tokens per batch
    k_lens: Optional[torch.Tensor] = None,  # (b,) valid key tokens per batch
    key_leftpad: Optional[torch.Tensor] = None,  # (b,)
    batch: int = 1,
) -> torch.Tensor:
    """Boolean (b, 1, sq, sk) tensor, True where key j is VISIBLE to query i.

    Geometry follows hopper/test_util.py:157-223: everything is aligned to the
    bottom-right corner, i.e. query i sits at absolute position
    ``i + k_len - q_len``; ``window_size=(l, r)`` keeps
    ``[pos - l, min(pos + r, k_len)]``; ``attention_chunk=c`` keeps the chunk
    ``[pos - pos % c, pos - pos % c + c)``; with ``key_leftpad`` column ``j``
    is logical key ``j - leftpad`` (columns left of the pad are never visible
    when a window is active; the key-length mask handles them otherwise).
    """
    if causal:
        window_size = (window_size[0], 0)
    wl, wr = window_size
    rows = torch.arange(seqlen_q, dtype=torch.long).view(1, 1, seqlen_q, 1)
    cols = torch.arange(seqlen_k, dtype=torch.long).view(1, 1, 1, seqlen_k).expand(batch, 1, 1, seqlen_k)
    if key_leftpad is not None:
        lp = key_leftpad.to(torch.long).view(batch, 1, 1, 1)
        cols = torch.where(cols >= lp, cols - lp, torch.full_like(cols, 2**32))
    kl = (
        torch.full((batch, 1, 1, 1), seqlen_k, dtype=torch.long)
        if k_lens is None
        else k_lens.to(torch.long).view(batch, 1, 1, 1)
    )
    ql = (
        torch.full((batch, 1, 1, 1), seqlen_q, dtype=torch.long)
        if q_lens is None
        else q_lens.to(torch.long).view(batch, 1, 1, 1)
    )
    pos = rows + kl - ql  # absolute position of each query row, (b,1,sq,1)
    visible = torch.ones(batch, 1, seqlen_q, seqlen_k, dtype=torch.bool)
    if wl >= 0 or wr >= 0:
        if wl < 0:
            hidden = cols > pos + wr
        else:
            # wr < 0 means "no right limit" (hopper/flash_api.cpp:152-153 turns it into
            # seqlen_q - 1).  The reference's attention_ref would compare against pos - 1 in
            # that never-tested corner; the kernel semantics are the ones restated here.
            right = torch.minimum(pos + wr, kl) if wr >= 0 else kl.expand_as(pos)
            hidden = (cols > right) | ((cols < pos - wl) & (cols >= sink_token_length))
        visible &= ~hidden
    if attention_chunk > 0:
        lo = pos - pos % attention_chunk
        visible &= ~((cols < lo) | (cols >= lo + attention_chunk))
    return visible


def _padding_mask(lens: Optional[torch.Tensor], seqlen: int, batch: int, leftpad=None):
    if lens is None and leftpad is None:
        return None
    idx = torch.arange(seqlen).view(1, seqlen)
    lo = torch.zeros(batch, 1, dtype=torch.long) if leftpad is None else leftpad.to(torch.long).view(batch, 1)
    n = torch.full((batch, 1), seqlen, dtype=torch.long) if lens is None else lens.to(torch.long).view(batch, 1)
    return (idx >= lo) & (idx < lo + n)


# --------------------------------------------------------------------------- attention
def attention_oracle(
    q: torch.Tensor,  # (b, sq, h, d)
    k: torch.Tensor,  # (b, sk, hk, d)
    v: torch.Tensor,  # (b, sk, hk, dv)
    *,
    q_lens: Optional[torch.Tensor] = None,
    k_lens: Optional[torch.Tensor] = None,
    key_leftpad: Optional[torch.Tensor] = None,
    causal: bool = False,
    window_size: Tuple[int, int] = (-1, -1),
    attention_chunk: int = 0,
    softcap: float = 0.0,
    softmax_scale: Optional[float] = None,
    q_descale: Optional[torch.Tensor] = None,  # (b, hk) fp32
    k_descale: Optional[torch.Tensor] = None,
    v_descale: Optional[torch.Tensor] = None,
    upcast: bool = True,
    reorder_ops: bool = False,
    intermediate_dtype: Optional[torch.dtype] = None,
    qv: Optional[torch.Tensor] = None,  # (b, sq, h, dv): scores += qv v^T (hopper/test_util.py:287-293)
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns ``(out (b,sq,h,dv) in q's dtype, lse (b,h,sq) fp32)``.

    ``k_lens`` counts valid keys starting at ``key_leftpad`` (0 if absent);
    rows beyond ``q_lens`` produce zeros.  Rows with no visible key give
    ``out = 0`` and ``lse = -inf`` (hopper/softmax.h:144-151).
    ``upcast=False, reorder_ops=True`` is the reference's "low precision"
    arm used to size the tolerance.
    """
    dtype_og = q.dtype
    b, sq, h, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    g = h // hk
    if upcast:
        q, k, v = q.float(), k.float(), v.float()
        qv = qv.float() if qv is not None else None
    if q_descale is not None:
        q = (q.float() * q_descale.repeat_interleave(g, dim=1).view(b, 1, h, 1)).to(q.dtype)
        qv = (qv.float() * q_descale.repeat_interleave(g, dim=1).view(b, 1, h, 1)).to(qv.dtype) if qv is not None else None
    if k_descale is not None:
        k = (k.float() * k_descale.view(b, 1, hk, 1)).to(k.dtype)
    if v_descale is not None:
        v = (v.float() * v_descale.view(b, 1, hk, 1)).to(v.dtype)
    k = k.repeat_interleave(g, dim=2)
    v = v.repeat_interleave(g, dim=2)
    # default scale with qv: 1 / sqrt(d + dv) (hopper/test_util.py:287)
    scale = 1.0 / math.sqrt(d if qv is None else d + v.shape[-1]) if softmax_scale is None else softmax_scale
    if reorder_ops:
        s = torch.einsum("bthd,bshd->bhts", q, k * scale)
    else:
        s = torch.einsum("bthd,bshd->bhts", q * scale, k)
    if qv is not None:
        s = s + torch.einsum("bthd,bshd->bhts", qv * scale, v)
    if softcap > 0:
        s = torch.tanh(s / softcap) * softcap
    kmask = _padding_mask(k_lens, sk, b, key_leftpad)
    qmask = _padding_mask(q_lens, sq, b)
    if kmask is not None:
        s = s.masked_fill(~kmask.view(b, 1, 1, sk), NEG_INF)
    has_shape_mask = causal or window_size[0] >= 0 or window_size[1] >= 0 or attention_chunk > 0
    if has_shape_mask:
        vis = visibility_mask(
            sq, sk, causal=causal, window_size=window_size, attention_chunk=attention_chunk,
            q_lens=q_lens, k_lens=k_lens, key_leftpad=key_leftpad, batch=b,
        )
        s = s.masked_fill(~vis, NEG_INF)
    sf = s.float()
    lse = torch.logsumexp(sf, dim=-1)  # (b,h,sq); -inf for empty rows
    p = torch.softmax(s, dim=-1).to(v.dtype)
    dead = torch.isneginf(sf).all(dim=-1, keepdim=True)
    p = p.masked_fill(dead, 0.0)
    if qmask is not None:
        p = p.masked_fill(~qmask.view(b, 1, sq, 1), 0.0)
        lse = lse.masked_fill(~qmask.view(b, 1, sq), NEG_INF)
    if intermediate_dtype is not None:
        p = p.to(intermediate_dtype).to(p.dtype)
    out = torch.einsum("bhts,bshd->bthd", p, v)
    if qmask is not None:
        out = out.masked_fill(~qmask.view(b, sq, 1, 1), 0.0)
    return out.to(dtype_og), lse


def attention_oracle_sliced(q, k, v, rows_per_call: int = 1024, **kw):
    """Exact oracle for shapes whose score matrix does not fit: one batch element and
    one block of query rows at a time.  For causal inputs the keys are cut at the last
    visible column (valid because the mask is bottom-right aligned)."""
    b, sq, h, _ = q.shape
    sk = k.shape[1]
    causal = kw.get("causal", False)
    assert kw.get("q_lens") is None and kw.get("k_lens") is None and kw.get("key_leftpad") is None
    outs, lses = [], []
    for bi in range(b):
        kwb = dict(kw)
        for name in ("q_descale", "k_descale", "v_descale"):
            if kwb.get(name) is not None:
                kwb[name] = kwb[name][bi : bi + 1]
        o_rows, l_rows = [], []
        for r0 in range(0, sq, rows_per_call):
            r1 = min(sq, r0 + rows_per_call)
            plain = kw.get("window_size", (-1, -1)) == (-1, -1) and kw.get("attention_chunk", 0) == 0
            if not plain:
                raise NotImplementedError("sliced oracle only supports plain / causal masks")
            if causal:
                # rows r0..r1 of an (sq, sk) problem == all rows of an (r1 - r0, kend) one
                kend = max(0, min(sk, r1 + sk - sq))
                if kend == 0:
                    o_rows.append(torch.zeros(1, r1 - r0, h, v.shape[-1], dtype=q.dtype))
                    l_rows.append(torch.full((1, h, r1 - r0), NEG_INF))
                    continue
                o, l = attention_oracle(q[bi : bi + 1, r0:r1], k[bi : bi + 1, :kend], v[bi : bi + 1, :kend], **kwb)
            else:
                o, l = attention_oracle(q[bi : bi + 1, r0:r1], k[bi : bi + 1], v[bi : bi + 1], **kwb)
            o_rows.append(o)
            l_rows.append(l)
        outs.append(torch.cat(o_rows, dim=1))
        lses.append(torch.cat(l_rows, dim=2))
    return torch.cat(outs, 0), torch.cat(lses, 0)


# --------------------------------------------------------------------------- combine
def combine_oracle(out_partial: torch.Tensor, lse_partial: torch.Tensor):
    """out_partial (splits, b, s, h, d) fp32, lse_partial (splits, b, s, h) fp32 (any strides).
    Returns out (b, s, h, d) fp32 and lse (b, s, h).  Splits whose LSE is -inf contribute 0."""
    lse = torch.logsumexp(lse_partial, dim=0)
    w = torch.exp(lse_partial - lse)
    w = torch.where(torch.isfinite(w), w, torch.zeros_like(w))
    out = (w.unsqueeze(-1) * out_partial).sum(0)
    return out, lse


# --------------------------------------------------------------------------- rotary
def rotary_oracle(x, cos, sin, seqlen_offsets=0, interleaved=False):
    """x (b, s, h, d); cos/sin (seqlen_ro, rotary_dim/2).  Token t of batch i is rotated with
    row ``seqlen_offsets[i] + t``; dims beyond rotary_dim pass through.  interleaved pairs
    (0,1),(2,3)..; otherwise (j, j + rotary_dim/2)  (hopper/rotary.h:19-77)."""
    b, s, h, d = x.shape
    rd = cos.shape[-1] * 2
    if isinstance(seqlen_offsets, int):
        seqlen_offsets = torch.full((b,), seqlen_offsets, dtype=torch.long)
    idx = seqlen_offsets.to(torch.long).view(b, 1) + torch.arange(s).view(1, s)
    c = cos.float()[idx].view(b, s, 1, rd // 2)
    sn = sin.float()[idx].view(b, s, 1, rd // 2)
    xr = x[..., :rd].float()
    if interleaved:
        x0, x1 = xr[..., 0::2], xr[..., 1::2]
        o = torch.stack((x0 * c - x1 * sn, x0 * sn + x1 * c), dim=-1).flatten(-2)
    else:
        x0, x1 = xr[..., : rd // 2], xr[..., rd // 2 :]
        o = torch.cat((x0 * c - x1 * sn, x0 * sn + x1 * c), dim=-1)
    return torch.cat((o.to(x.dtype), x[..., rd:]), dim=-1)


# --------------------------------------------------------------------------- paged cache
def paged_kvcache(seqlen_k, page_size, batch, hk, d, dv, dtype, generator=None, pool_factor=3):
    """Random paged cache + the dense view it encodes (hopper/test_flash_attn.py:992-1015)."""
    per_seq = math.ceil(seqlen_k / page_size)
    num_pages = per_seq * batch * pool_factor
    kp = torch.randn(num_pages, page_size, hk, d, generator=generator).to(dtype)
    vp = torch.randn(num_pages, page_size, hk, dv, generator=generator).to(dtype)
    table = torch.randperm(num_pages, generator=generator).to(torch.int32).view(batch, -1)
    kd = kp[table.flatten().long()].view(batch, -1, hk, d)[:, :seqlen_k]
    vd = vp[table.flatten().long()].view(batch, -1, hk, dv)[:, :seqlen_k]
    return kd, vd, table, kp, vp


# --------------------------------------------------------------------------- tolerance
def fwd_tolerance(out_ref: torch.Tensor, out_lowprec: torch.Tensor, rtol: float = 2.0) -> float:
    """max|out - out_ref| must be <= this (hopper/test_flash_attn.py:193-223)."""
    ref = out_ref.float()
    atol = 2 * (ref + 0.3 - 0.3 - ref).abs().max().item()
    return rtol * (out_lowprec.float() - ref).abs().max().item() + atol


def flops(batch, nheads, seqlen_q, seqlen_k, headdim, headdim_v, causal=False):
    """Algorithmic FLOPs, causal-adjusted (hopper/benchmark_attn.py:62-73)."""
    avg_k = (max(0, seqlen_k - seqlen_q) + seqlen_k) / 2 if causal else seqlen_k
    return batch * nheads * 2 * seqlen_q * avg_k * (headdim + headdim_v)
