"""GPU parity tests: the CUDA path (flash_attn_interface -> torch.ops.flash_attn_3 -> libfa100.so C ABI) against the
CPU oracle on the same seeded inputs.  Shapes and tolerance rule follow hopper/test_flash_attn.py:53-286 (output),
:290-598 (varlen), :601-989 (kvcache), :1043-1102 (bitwise repeatability).

Tolerances (written here as north_star requires):
  bf16/fp16: max|out - ref_fp32| <= 2 * max|ref_lowprec - ref_fp32| + fwd_atol   (3x with softcap)
  fp8:       same rule with the e4m3-intermediate low-precision arm and multiplier 4
  LSE:       |lse - lse_ref| <= 1e-3 * max(1, |lse_ref|), and -inf exactly where the oracle has -inf
"""
import math

import os

import pytest
import torch

import attention_oracle as A

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _iface():
    import flash_attn_interface as F

    return F


def _check(out, lse, ref, low, lse_ref, rtol=2.0):
    import fa100_native as N

    torch.cuda.synchronize()
    hang = N.hang_info()
    assert hang[0] == 0, f"kernel watchdog fired: {[hex(x) for x in hang]}"
    assert torch.isfinite(out.float()).all()
    err = (out.float().cpu() - ref.float()).abs().max().item()
    tol = A.fwd_tolerance(ref, low, rtol)
    assert err <= tol, f"max err {err:.3e} > tol {tol:.3e}"
    if lse is not None:
        lse = lse.cpu()
        assert torch.equal(torch.isneginf(lse), torch.isneginf(lse_ref))
        fin = torch.isfinite(lse_ref)
        rel = ((lse[fin] - lse_ref[fin]).abs() / lse_ref[fin].abs().clamp_min(1.0)).max().item() if fin.any() else 0.0
        assert rel <= 1e-3, f"lse rel err {rel:.3e}"


def _qkv(b, sq, sk, h, hk, d, dtype, seed=0, dv=None):
    g = torch.Generator().manual_seed(seed)
    dv = dv or d
    return (torch.randn(b, sq, h, d, generator=g).to(dtype), torch.randn(b, sk, hk, d, generator=g).to(dtype),
            torch.randn(b, sk, hk, dv, generator=g).to(dtype))


SHAPES = [(1, 1), (64, 128), (128, 128), (256, 256), (113, 203), (128, 217), (113, 211), (108, 256), (256, 512), (239, 1),
          (384, 256), (640, 128), (512, 256), (1024, 1024), (1023, 1024), (799, 3), (1, 1024)]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("mha_type", ["mha", "mqa", "gqa"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128])
@pytest.mark.parametrize("sq,sk", SHAPES)
def test_output(sq, sk, d, causal, mha_type, dtype):
    F = _iface()
    h = 6
    hk = {"mha": 6, "mqa": 1, "gqa": 2}[mha_type]
    q, k, v = _qkv(3, sq, sk, h, hk, d, dtype)
    ref, lse_ref = A.attention_oracle(q, k, v, causal=causal)
    low, _ = A.attention_oracle(q, k, v, causal=causal, upcast=False, reorder_ops=True)
    for pack_gqa, num_splits in ((False, 1), (True, 1), (False, 3), (True, 3)):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), causal=causal, pack_gqa=pack_gqa, num_splits=num_splits,
                                     return_attn_probs=True)
        _check(out, lse, ref, low, lse_ref)


@pytest.mark.parametrize("kw", [dict(window_size=(100, 50)), dict(window_size=(300, 0)), dict(window_size=(0, 0)),
                                dict(window_size=(-1, 30)), dict(attention_chunk=128), dict(attention_chunk=77, window_size=(200, 10)),
                                dict(softcap=15.0), dict(softcap=15.0, causal=True), dict(softmax_scale=0.3, causal=True)])
@pytest.mark.parametrize("sq,sk", [(500, 700), (700, 500), (64, 1500), (1, 333)])
def test_masks_and_softcap(sq, sk, kw):
    F = _iface()
    q, k, v = _qkv(2, sq, sk, 4, 2, 128, torch.bfloat16, seed=1)
    okw = dict(kw)
    ref, lse_ref = A.attention_oracle(q, k, v, **okw)
    low, _ = A.attention_oracle(q, k, v, upcast=False, reorder_ops=True, **okw)
    for num_splits in (1, 2):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), return_attn_probs=True, num_splits=num_splits, **kw)
        _check(out, lse, ref, low, lse_ref, rtol=3.0 if kw.get("softcap") else 2.0)


@pytest.mark.parametrize("d,dv", [(96, 96), (80, 80), (128, 64), (64, 128), (40, 40)])
def test_odd_head_dims(d, dv):
    F = _iface()
    q, k, v = _qkv(2, 300, 411, 4, 2, d, torch.bfloat16, seed=2, dv=dv)
    ref, lse_ref = A.attention_oracle(q, k, v, causal=True)
    low, _ = A.attention_oracle(q, k, v, causal=True, upcast=False, reorder_ops=True)
    out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), causal=True, return_attn_probs=True)
    _check(out, lse, ref, low, lse_ref)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d,dv", [(256, 256), (192, 192), (192, 128), (160, 160), (136, 136), (64, 256), (256, 64)])
@pytest.mark.parametrize("sq,sk", [(1, 1), (128, 128), (113, 203), (300, 411), (1, 1024), (640, 128), (1023, 1024)])
def test_large_head_dims(sq, sk, d, dv, causal, dtype):
    """head dims in (128, 256] (hopper/flash_api.cpp:710, tile_size_fwd_sm90 hdim 192 / 256 rows): the one-Q-tile-per-CTA
    instance, every scheduling mode."""
    F = _iface()
    q, k, v = _qkv(2, sq, sk, 4, 2, d, dtype, seed=3, dv=dv)
    ref, lse_ref = A.attention_oracle(q, k, v, causal=causal)
    low, _ = A.attention_oracle(q, k, v, causal=causal, upcast=False, reorder_ops=True)
    for pack_gqa, num_splits in ((False, 1), (True, 1), (False, 3), (True, 2)):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), causal=causal, pack_gqa=pack_gqa, num_splits=num_splits,
                                     return_attn_probs=True)
        _check(out, lse, ref, low, lse_ref)


def test_large_head_dim_kvcache_varlen():
    """hdim 256 through the paged KV cache (split chosen by the plan) and the varlen entry point."""
    F = _iface()
    b, h, hk, d, ps, cap = 3, 4, 2, 256, 64, 1024
    g = torch.Generator().manual_seed(11)
    q = torch.randn(b, 2, h, d, generator=g).to(torch.bfloat16)
    kc = torch.randn(b, cap, hk, d, generator=g).to(torch.bfloat16)
    vc = torch.randn(b, cap, hk, d, generator=g).to(torch.bfloat16)
    lens = torch.tensor([1024, 517, 1], dtype=torch.int32)
    ref, lse_ref = A.attention_oracle(q, kc, vc, causal=True, k_lens=lens)
    low, _ = A.attention_oracle(q, kc, vc, causal=True, k_lens=lens, upcast=False, reorder_ops=True)
    npg = cap // ps
    perm = torch.randperm(b * npg, generator=g).to(torch.int32)
    table = perm.view(b, npg)
    kp = torch.empty(b * npg, ps, hk, d, dtype=torch.bfloat16)
    vp = torch.empty_like(kp)
    kp[table.flatten().long()] = kc.view(b * npg, ps, hk, d)
    vp[table.flatten().long()] = vc.view(b * npg, ps, hk, d)
    for kk, vv, tb in ((kc, vc, None), (kp, vp, table)):
        out, lse = F.flash_attn_with_kvcache(q.to(DEV), kk.to(DEV), vv.to(DEV), cache_seqlens=lens.to(DEV), causal=True,
                                             page_table=None if tb is None else tb.to(DEV), return_softmax_lse=True)[:2]
        _check(out, lse, ref, low, lse_ref)
    lens_q, lens_k = [5, 200, 131], [64, 333, 131]
    qs = [torch.randn(n, h, d, generator=g).to(torch.bfloat16) for n in lens_q]
    ks = [torch.randn(n, hk, d, generator=g).to(torch.bfloat16) for n in lens_k]
    vs = [torch.randn(n, hk, d, generator=g).to(torch.bfloat16) for n in lens_k]
    cu = lambda ls: torch.tensor([0] + list(torch.tensor(ls).cumsum(0)), dtype=torch.int32).to(DEV)
    out, lse = F.flash_attn_varlen_func(torch.cat(qs).to(DEV), torch.cat(ks).to(DEV), torch.cat(vs).to(DEV), cu(lens_q), cu(lens_k),
                                        max(lens_q), max(lens_k), causal=True, return_attn_probs=True)
    off = 0
    for qi, ki, vi in zip(qs, ks, vs):
        ref, lse_ref = A.attention_oracle(qi[None], ki[None], vi[None], causal=True)
        low, _ = A.attention_oracle(qi[None], ki[None], vi[None], causal=True, upcast=False, reorder_ops=True)
        n = qi.shape[0]
        _check(out[off:off + n][None], lse[:, off:off + n][None], ref, low, lse_ref)
        off += n


def test_noncontiguous_inputs():
    """only last-dim contiguity is required (hopper/flash_api.cpp:726-728)."""
    F = _iface()
    q, k, v = _qkv(2, 200, 200, 4, 4, 128, torch.bfloat16, seed=3)
    qkv = torch.stack([q, k, v], dim=2).to(DEV)  # (b, s, 3, h, d): packed QKV views
    out = F.flash_attn_qkvpacked_func(qkv, causal=True)
    ref, _ = A.attention_oracle(q, k, v, causal=True)
    low, _ = A.attention_oracle(q, k, v, causal=True, upcast=False, reorder_ops=True)
    _check(out, None, ref, low, None)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("lens_q,lens_k", [([100, 257, 31], [100, 300, 77]), ([128, 5, 300, 1], [128, 0, 511, 1000]),
                                           ([1, 1, 1, 1, 1], [7, 130, 250, 1, 600])])
def test_varlen(lens_q, lens_k, causal):
    F = _iface()
    g = torch.Generator().manual_seed(4)
    h, hk, d = 4, 2, 128
    cq = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), dtype=torch.int32)
    ck = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), dtype=torch.int32)
    q = torch.randn(int(cq[-1]), h, d, generator=g).to(torch.bfloat16)
    k = torch.randn(int(ck[-1]), hk, d, generator=g).to(torch.bfloat16)
    v = torch.randn(int(ck[-1]), hk, d, generator=g).to(torch.bfloat16)
    refs, lows, lses = [], [], []
    for i in range(len(lens_q)):
        qi, ki, vi = q[cq[i]:cq[i + 1]][None], k[ck[i]:ck[i + 1]][None], v[ck[i]:ck[i + 1]][None]
        if lens_k[i] == 0:
            refs.append(torch.zeros_like(qi[0])); lows.append(torch.zeros_like(qi[0]))
            lses.append(torch.full((h, lens_q[i]), float("-inf")))
            continue
        r, l = A.attention_oracle(qi, ki, vi, causal=causal)
        lo, _ = A.attention_oracle(qi, ki, vi, causal=causal, upcast=False, reorder_ops=True)
        refs.append(r[0]); lows.append(lo[0]); lses.append(l[0])
    for pack_gqa, num_splits in ((False, 1), (True, 2)):
        out, lse = F.flash_attn_varlen_func(q.to(DEV), k.to(DEV), v.to(DEV), cq.to(DEV), ck.to(DEV), max(lens_q), max(lens_k),
                                            causal=causal, pack_gqa=pack_gqa, num_splits=num_splits, return_attn_probs=True)
        _check(out, lse, torch.cat(refs), torch.cat(lows), torch.cat(lses, dim=1))


def test_varlen_seqused():
    """seqused_q / seqused_k shorter than the cu_seqlens slots (hopper/test_flash_attn.py add_unused_qkv)."""
    F = _iface()
    g = torch.Generator().manual_seed(5)
    b, sq, sk, h, hk, d = 3, 160, 300, 4, 1, 128
    q, k, v = _qkv(b, sq, sk, h, hk, d, torch.bfloat16, seed=5)
    uq = torch.tensor([160, 17, 100], dtype=torch.int32)
    uk = torch.tensor([300, 129, 5], dtype=torch.int32)
    cq = torch.arange(0, (b + 1) * sq, sq, dtype=torch.int32)
    ck = torch.arange(0, (b + 1) * sk, sk, dtype=torch.int32)
    out, lse = F.flash_attn_varlen_func(q.reshape(-1, h, d).to(DEV), k.reshape(-1, hk, d).to(DEV), v.reshape(-1, hk, d).to(DEV),
                                        cq.to(DEV), ck.to(DEV), sq, sk, seqused_q=uq.to(DEV), seqused_k=uk.to(DEV), causal=True,
                                        return_attn_probs=True)
    ref, lse_ref = A.attention_oracle(q, k, v, q_lens=uq, k_lens=uk, causal=True)
    low, _ = A.attention_oracle(q, k, v, q_lens=uq, k_lens=uk, causal=True, upcast=False, reorder_ops=True)
    valid = (torch.arange(sq).view(1, -1) < uq.view(-1, 1))
    out = torch.where(valid.view(b, sq, 1, 1), out.view(b, sq, h, d).cpu(), torch.zeros((), dtype=out.dtype))  # rows past seqused_q are not written
    lse = lse.view(h, b, sq).permute(1, 0, 2).cpu().masked_fill(~valid.view(b, 1, sq), float("-inf"))
    _check(out, lse, ref, low, lse_ref)


@pytest.mark.parametrize("page_size", [None, 16, 64, 128, 256])
@pytest.mark.parametrize("sq,sk", [(1, 128), (1, 339), (3, 1024), (64, 800), (16, 3000), (128, 128)])
@pytest.mark.parametrize("causal", [False, True])
def test_kvcache(sq, sk, page_size, causal):
    F = _iface()
    g = torch.Generator().manual_seed(6)
    b, h, hk, d = 4, 8, 2, 128
    q = torch.randn(b, sq, h, d, generator=g).to(torch.bfloat16)
    lens = torch.randint(max(1, sk // 2), sk + 1, (b,), generator=g, dtype=torch.int32)
    if causal:
        lens = lens.clamp_min(sq)
    if page_size is None:
        kd = torch.randn(b, sk, hk, d, generator=g).to(torch.bfloat16)
        vd = torch.randn(b, sk, hk, d, generator=g).to(torch.bfloat16)
        kin, vin, table = kd, vd, None
    else:
        kd, vd, table, kin, vin = A.paged_kvcache(sk, page_size, b, hk, d, d, torch.bfloat16, generator=g)
    ref, lse_ref = A.attention_oracle(q, kd, vd, k_lens=lens, causal=causal)
    low, _ = A.attention_oracle(q, kd, vd, k_lens=lens, causal=causal, upcast=False, reorder_ops=True)
    for num_splits in (0, 1, 3):
        out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kin.to(DEV), vin.to(DEV), cache_seqlens=lens.to(DEV),
                                                 page_table=None if table is None else table.to(DEV), causal=causal,
                                                 num_splits=num_splits, return_softmax_lse=True)
        _check(out, lse, ref, low, lse_ref)


@pytest.mark.parametrize("page_size", [1, 3, 8, 24, 48, 96, 384])
@pytest.mark.parametrize("sq,sk", [(1, 339), (16, 1000)])
def test_kvcache_any_page_size(sq, sk, page_size):
    """Page sizes that neither divide nor are a multiple of the 128-key block (the reference serves them with its
    cp.async gather path, hopper/paged_kv.h:233-283; here gcd(page, 128) rows per TMA box)."""
    F = _iface()
    g = torch.Generator().manual_seed(60 + page_size)
    b, h, hk, d = 3, 8, 2, 128
    q = torch.randn(b, sq, h, d, generator=g).to(torch.bfloat16)
    lens = torch.randint(max(sq, sk // 2), sk + 1, (b,), generator=g, dtype=torch.int32)
    kd, vd, table, kin, vin = A.paged_kvcache(sk, page_size, b, hk, d, d, torch.bfloat16, generator=g)
    ref, lse_ref = A.attention_oracle(q, kd, vd, k_lens=lens, causal=True)
    low, _ = A.attention_oracle(q, kd, vd, k_lens=lens, causal=True, upcast=False, reorder_ops=True)
    for num_splits in (0, 3):
        out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kin.to(DEV), vin.to(DEV), cache_seqlens=lens.to(DEV),
                                                 page_table=table.to(DEV), causal=True, num_splits=num_splits,
                                                 return_softmax_lse=True)
        _check(out, lse, ref, low, lse_ref)


def test_kvcache_batch_idx_and_leftpad():
    F = _iface()
    g = torch.Generator().manual_seed(7)
    b, bc, sq, sk, h, hk, d = 3, 7, 2, 600, 4, 2, 128
    q = torch.randn(b, sq, h, d, generator=g).to(torch.bfloat16)
    kc = torch.randn(bc, sk, hk, d, generator=g).to(torch.bfloat16)
    vc = torch.randn(bc, sk, hk, d, generator=g).to(torch.bfloat16)
    idx = torch.tensor([5, 0, 3], dtype=torch.int32)
    lens = torch.tensor([600, 257, 400], dtype=torch.int32)   # cache_seqlens counts from the start of the buffer
    pad = torch.tensor([0, 100, 37], dtype=torch.int32)
    out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kc.to(DEV), vc.to(DEV), cache_seqlens=lens.to(DEV), cache_batch_idx=idx.to(DEV),
                                             cache_leftpad=pad.to(DEV), causal=True, return_softmax_lse=True)
    kd, vd = kc[idx.long()], vc[idx.long()]
    ref, lse_ref = A.attention_oracle(q, kd, vd, k_lens=lens - pad, key_leftpad=pad, causal=True)
    low, _ = A.attention_oracle(q, kd, vd, k_lens=lens - pad, key_leftpad=pad, causal=True, upcast=False, reorder_ops=True)
    _check(out, lse, ref, low, lse_ref)


@pytest.mark.parametrize("rotary", [None, "interleaved", "halves", "halves_seqlens", "partial"])
@pytest.mark.parametrize("layout", ["dense", "batch_idx", "paged64", "paged256", "varlen_new"])
@pytest.mark.parametrize("mask", ["none", "causal", "local"])
@pytest.mark.parametrize("sq,s_new", [(1, 1), (5, 5), (64, 37), (3, 130)])
def test_kvcache_append_rotary(sq, s_new, mask, layout, rotary):
    """k / v appended in place at cache_seqlens, rotary on k_new and on a copy of q (hopper/test_flash_attn.py:601-989):
    the cache must hold exactly the new rows (bit-exact without rotary, 1e-3 with) and nothing else may change."""
    if sq > s_new:
        pytest.skip("seqlen_q <= seqlen_new in the reference test")
    F = _iface()
    g = torch.Generator().manual_seed(11)
    b, h, hk, d, cap = 3, 4, 2, 128, 512
    dt = torch.bfloat16
    q = torch.randn(b, sq, h, d, generator=g).to(dt)
    kn = torch.randn(b, s_new, hk, d, generator=g).to(dt)
    vn = torch.randn(b, s_new, hk, d, generator=g).to(dt)
    lens = torch.randint(0, cap - s_new + 1, (b,), generator=g, dtype=torch.int32)
    new_lens = torch.full((b,), s_new, dtype=torch.int32)
    if layout == "varlen_new":
        new_lens = torch.randint(1, s_new + 1, (b,), generator=g, dtype=torch.int32)
        new_lens[0] = s_new
    causal, window = mask == "causal", ((40, 10) if mask == "local" else (-1, -1))
    if causal or mask == "local":  # keep sq <= the shortest final key length so no query row is fully masked
        lens = lens.clamp_min(sq)
    nb = 5 if layout == "batch_idx" else b
    idx = torch.tensor([4, 1, 2], dtype=torch.int32) if layout == "batch_idx" else None
    table = None
    if layout.startswith("paged"):
        ps = int(layout[5:])
        kd, vd, table, kc, vc = A.paged_kvcache(cap, ps, b, hk, d, d, dt, generator=g)
        kd, vd = kd.clone(), vd.clone()
    else:
        kc = torch.randn(nb, cap, hk, d, generator=g).to(dt)
        vc = torch.randn(nb, cap, hk, d, generator=g).to(dt)
        sel = idx.long() if idx is not None else torch.arange(b)
        kd, vd = kc[sel].clone(), vc[sel].clone()
    cos = sin = rot_lens = None
    q_ref, kn_ref = q, kn
    if rotary is not None:
        rd = 64 if rotary == "partial" else d
        ang = torch.rand(cap + 64, rd // 2, generator=g) * 2 * math.pi
        cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)
        inter = rotary == "interleaved"
        rot_lens = (lens // 2) if rotary == "halves_seqlens" else None
        base = rot_lens if rot_lens is not None else lens
        kn_ref = A.rotary_oracle(kn, cos, sin, base, interleaved=inter)
        if causal or mask == "local":
            q_ref = A.rotary_oracle(q, cos, sin, base, interleaved=inter)
        else:  # every query row sits at position cache_seqlens
            q_ref = A.rotary_oracle(q.reshape(b, 1, sq * h, d), cos, sin, base, interleaved=inter).reshape(b, sq, h, d)
    for i in range(b):  # the dense view after the append
        n = int(new_lens[i])
        kd[i, int(lens[i]):int(lens[i]) + n] = kn_ref[i, :n]
        vd[i, int(lens[i]):int(lens[i]) + n] = vn[i, :n]
    tot = lens + new_lens
    kw = dict(causal=causal, window_size=window)
    ref, lse_ref = A.attention_oracle(q_ref, kd, vd, k_lens=tot, **kw)
    low, _ = A.attention_oracle(q_ref, kd, vd, k_lens=tot, upcast=False, reorder_ops=True, **kw)
    kc_dev, vc_dev = kc.to(DEV), vc.to(DEV)
    k_before, v_before = kc_dev.clone(), vc_dev.clone()
    if layout == "varlen_new":
        cu_new = torch.nn.functional.pad(new_lens.cumsum(0), (1, 0)).to(torch.int32)
        kn_in = torch.cat([kn[i, :int(new_lens[i])] for i in range(b)]).to(DEV)
        vn_in = torch.cat([vn[i, :int(new_lens[i])] for i in range(b)]).to(DEV)
        extra = dict(cu_seqlens_k_new=cu_new.to(DEV))
    else:
        kn_in, vn_in, extra = kn.to(DEV), vn.to(DEV), {}
    q_dev = q.to(DEV)
    q_copy = q_dev.clone()
    out, lse, *_ = F.flash_attn_with_kvcache(
        q_dev, kc_dev, vc_dev, k=kn_in, v=vn_in, rotary_cos=None if cos is None else cos.to(DEV),
        rotary_sin=None if sin is None else sin.to(DEV), cache_seqlens=lens.to(DEV),
        cache_batch_idx=None if idx is None else idx.to(DEV), page_table=None if table is None else table.to(DEV),
        rotary_seqlens=None if rot_lens is None else rot_lens.to(DEV), rotary_interleaved=(rotary == "interleaved"),
        return_softmax_lse=True, **kw, **extra)
    _check(out, lse, ref, low, lse_ref)
    assert torch.equal(q_dev, q_copy), "q must not be modified"
    # cache contents: dense view of what the device now holds
    if table is not None:
        kd_dev = kc_dev.cpu()[table.flatten().long()].view(b, -1, hk, d)[:, :cap]
        vd_dev = vc_dev.cpu()[table.flatten().long()].view(b, -1, hk, d)[:, :cap]
    else:
        sel = idx.long() if idx is not None else torch.arange(b)
        kd_dev, vd_dev = kc_dev.cpu()[sel], vc_dev.cpu()[sel]
    assert torch.equal(vd_dev, vd)
    if rotary is None:
        assert torch.equal(kd_dev, kd)
    else:
        assert torch.allclose(kd_dev.float(), kd.float(), rtol=1e-3, atol=1e-3)
    # rows outside the appended ranges are untouched (also the cache entries no batch element points at)
    changed_k = (kc_dev != k_before).flatten(1).any(1).sum().item() if table is None else None
    if table is None:
        assert changed_k <= b
        untouched = [i for i in range(nb) if idx is None or i not in idx.tolist()] if idx is not None else []
        for i in untouched:
            assert torch.equal(kc_dev[i], k_before[i]) and torch.equal(vc_dev[i], v_before[i])


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d,dv", [(128, 128), (64, 64), (96, 96), (256, 256), (192, 128)])
@pytest.mark.parametrize("sq,sk", [(128, 128), (700, 700), (113, 203), (1, 900)])
def test_fp8(sq, sk, d, dv, causal):
    """e4m3 inputs, fp32 (b, h_k) descales, bf16 output (hopper/test_flash_attn.py:118-149, multipliers :986-989)."""
    F = _iface()
    g = torch.Generator().manual_seed(8)
    b, h, hk = 2, 4, 2
    q = torch.randn(b, sq, h, d, generator=g).to(torch.float8_e4m3fn)
    k = torch.randn(b, sk, hk, d, generator=g).to(torch.float8_e4m3fn)
    v = torch.randn(b, sk, hk, dv, generator=g).to(torch.float8_e4m3fn)
    ds = [torch.rand(b, hk, generator=g) * 2 for _ in range(3)]
    qr, kr, vr = (x.to(torch.bfloat16) for x in (q, k, v))
    kw = dict(causal=causal, q_descale=ds[0], k_descale=ds[1], v_descale=ds[2])
    ref, lse_ref = A.attention_oracle(qr, kr, vr, **kw)
    low, _ = A.attention_oracle(qr, kr, vr, upcast=False, reorder_ops=True, intermediate_dtype=torch.float8_e4m3fn, **kw)
    out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), causal=causal, q_descale=ds[0].to(DEV), k_descale=ds[1].to(DEV),
                                 v_descale=ds[2].to(DEV), return_attn_probs=True)
    assert out.dtype == torch.bfloat16
    _check(out, lse, ref, low, lse_ref, rtol=4.0)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float8_e4m3fn])
@pytest.mark.parametrize("step", [0.5, 2.0, 6.0])
def test_growing_row_max(step, dtype):
    """Scores that keep growing in the order the kernel walks the key blocks (last block first): every block raises the row
    max by `step` nats, so the lazy rescale (stale max within 2^8, or 2^2.75 for e4m3 P) and the accumulator rescale in
    tensor memory are exercised at every step; random data almost never triggers them."""
    F = _iface()
    g = torch.Generator().manual_seed(21)
    b, h, sq, sk, d = 1, 2, 160, 1280, 128
    q = torch.randn(b, sq, h, d, generator=g)
    k = torch.randn(b, sk, h, d, generator=g) * 0.1
    v = torch.randn(b, sk, h, d, generator=g)
    q[..., 0] = 4.0                                            # s = q . k / sqrt(d): the first component carries the ramp
    ramp = (sk // 128 - 1 - torch.arange(sk) // 128).float() * step * (d ** 0.5) / 4.0
    k[..., 0] = ramp[None, :, None]
    lo = torch.float8_e4m3fn if dtype == torch.float8_e4m3fn else dtype
    q, k, v = (x.to(lo) for x in (q, k, v))
    qr, kr, vr = (x.to(torch.bfloat16) for x in (q, k, v)) if dtype == torch.float8_e4m3fn else (q, k, v)
    ref, lse_ref = A.attention_oracle(qr, kr, vr)
    kw = dict(intermediate_dtype=torch.float8_e4m3fn) if dtype == torch.float8_e4m3fn else {}
    low, _ = A.attention_oracle(qr, kr, vr, upcast=False, reorder_ops=True, **kw)
    for num_splits in (1, 3):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), num_splits=num_splits, return_attn_probs=True)
        _check(out, lse, ref, low, lse_ref, rtol=4.0 if dtype == torch.float8_e4m3fn else 2.0)


@pytest.mark.parametrize("num_splits", [1, 2, 3, 5, 17, 32, 55, 97, 133])
@pytest.mark.parametrize("seqlen", [1, 3, 113, 640])
@pytest.mark.parametrize("d,dtype", [(64, torch.float32), (128, torch.bfloat16), (96, torch.float16), (256, torch.float32)])
def test_combine(num_splits, seqlen, d, dtype):
    """hopper/test_flash_attn.py:1117-1156: non-contiguous inputs, -inf splits, lse to 1e-5."""
    F = _iface()
    torch.manual_seed(1)
    b, h = 5, 16
    op = torch.randn(num_splits * 2, b, h, seqlen, d, device=DEV).transpose(2, 3)[:num_splits]
    lp = torch.randn(num_splits, b, h * 2, seqlen, device=DEV).transpose(-1, -2)[:, :, :, :h]
    lp[num_splits // 2:, : b // 3] = float("-inf")
    out, lse = F.flash_attn_combine(op, lp, out_dtype=dtype)
    oref, lref = A.combine_oracle(op.cpu(), lp.cpu())
    assert torch.allclose(lse.cpu(), lref, atol=1e-5, rtol=1e-5)
    opt = oref.to(dtype)
    err = (out.cpu().float() - oref).abs().max().item()
    assert err <= 2 * (opt.float() - oref).abs().max().item() or torch.allclose(out.cpu().float(), opt.float(), atol=1e-5, rtol=1e-5)


def test_bitwise_repeatable():
    """hopper/test_flash_attn.py:1043-1102: the forward must be bitwise deterministic run to run."""
    F = _iface()
    q, k, v = (x.to(DEV) for x in _qkv(2, 1024, 1024, 8, 2, 128, torch.bfloat16, seed=9))
    for kw in (dict(causal=True), dict(causal=True, num_splits=4), dict(causal=False, pack_gqa=True)):
        o0, l0 = F.flash_attn_func(q, k, v, return_attn_probs=True, **kw)
        for _ in range(50):
            o, l = F.flash_attn_func(q, k, v, return_attn_probs=True, **kw)
            assert torch.equal(o, o0) and torch.equal(l, l0)


def test_split_combine_identity_full_size():
    """size-independent property at a large size: split-KV partials merged by the combine kernel == unsplit result."""
    F = _iface()
    q, k, v = (x.to(DEV) for x in _qkv(1, 2048, 8192, 8, 2, 128, torch.bfloat16, seed=10))
    o1, l1 = F.flash_attn_func(q, k, v, causal=True, num_splits=1, return_attn_probs=True)
    o8, l8 = F.flash_attn_func(q, k, v, causal=True, num_splits=8, return_attn_probs=True)
    assert (o1.float() - o8.float()).abs().max().item() <= 2e-2
    assert torch.allclose(l1, l8, atol=1e-4, rtol=1e-4)


def _gpu_randn(shape, seed, dtype=torch.bfloat16):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return torch.randn(*shape, generator=g, device=DEV, dtype=torch.bfloat16).to(dtype)


def _slice_check(out, lse, q, k, v, bi, g, G, rows, causal, rtol=2.0, **kw):
    """Oracle on one (batch, kv head) and one block of query rows of a problem too large to check whole; for causal
    inputs rows r0..r1 of an (s, s) problem equal all rows of an (r1 - r0, r1) one (bottom-right aligned mask)."""
    r0, r1 = rows
    kend = r1 if causal else k.shape[1]
    hs = slice(g * G, (g + 1) * G)
    qs, ks, vs = q[bi:bi + 1, r0:r1, hs].cpu(), k[bi:bi + 1, :kend, g:g + 1].cpu(), v[bi:bi + 1, :kend, g:g + 1].cpu()
    kwb = {n: t[bi:bi + 1, g:g + 1].cpu() for n, t in kw.items()}
    if qs.dtype == torch.float8_e4m3fn:
        qs, ks, vs = (x.to(torch.bfloat16) for x in (qs, ks, vs))
        low_kw = dict(intermediate_dtype=torch.float8_e4m3fn)
    else:
        low_kw = {}
    ref, lse_ref = A.attention_oracle(qs, ks, vs, causal=causal, **kwb)
    low, _ = A.attention_oracle(qs, ks, vs, causal=causal, upcast=False, reorder_ops=True, **low_kw, **kwb)
    _check(out[bi:bi + 1, r0:r1, hs], lse[bi:bi + 1, hs, r0:r1], ref, low, lse_ref, rtol=rtol)


def test_c2_full_size():
    """BASELINE config C2 (bf16 b8 32q/8kv s8192 d128 causal) at full size: oracle on slices, and the size-independent
    properties the domain offers -- batch permutation equivariance and exact linearity in V (both bitwise)."""
    F = _iface()
    b, s, h, hk, d = 8, 8192, 32, 8, 128
    q, k, v = _gpu_randn((b, s, h, d), 21), _gpu_randn((b, s, hk, d), 22), _gpu_randn((b, s, hk, d), 23)
    out, lse = F.flash_attn_func(q, k, v, causal=True, return_attn_probs=True)
    for bi, g, rows in ((0, 0, (0, 256)), (3, 5, (4000, 4256)), (7, 7, (7936, 8192)), (5, 2, (127, 130))):
        _slice_check(out, lse, q, k, v, bi, g, h // hk, rows, True)
    perm = torch.tensor([3, 0, 7, 1, 6, 2, 5, 4], device=DEV)
    out_p, lse_p = F.flash_attn_func(q[perm], k[perm], v[perm], causal=True, return_attn_probs=True)
    assert torch.equal(out_p, out[perm]) and torch.equal(lse_p, lse[perm])
    out_2, lse_2 = F.flash_attn_func(q, k, v * 2, causal=True, return_attn_probs=True)
    assert torch.equal(out_2, out * 2) and torch.equal(lse_2, lse)


@pytest.mark.parametrize("causal", [True, False])
def test_c3_full_size(causal):
    """BASELINE config C3 (e4m3 b4 h16 s16384 d128, per-tensor descales) at full size: oracle on slices, V linearity."""
    F = _iface()
    b, s, h, d = 4, 16384, 16, 128
    q, k, v = (_gpu_randn((b, s, h, d), 31 + i, torch.float8_e4m3fn) for i in range(3))
    ds = {n: torch.full((b, h), x, dtype=torch.float32, device=DEV) for n, x in (("q_descale", 0.5), ("k_descale", 0.75), ("v_descale", 1.25))}
    out, lse = F.flash_attn_func(q, k, v, causal=causal, return_attn_probs=True, **ds)
    for bi, g, rows in ((0, 0, (0, 128)), (2, 9, (8100, 8228)), (3, 15, (16256, 16384))):
        _slice_check(out, lse, q, k, v, bi, g, 1, rows, causal, rtol=4.0, **ds)
    ds2 = dict(ds, v_descale=ds["v_descale"] * 2)
    out_2, lse_2 = F.flash_attn_func(q, k, v, causal=causal, return_attn_probs=True, **ds2)
    assert torch.equal(out_2, out * 2) and torch.equal(lse_2, lse)


@pytest.mark.parametrize("ragged", [False, True])
def test_c4_full_size(ragged):
    """BASELINE config C4 (decode b64 32q/8kv sq1 sk32768 d128, page 256, split-KV + combine) at full size: oracle on
    whole sequences of a few batch elements; auto split equals forced 1 / 4 splits within the combine tolerance."""
    F = _iface()
    b, sk, h, hk, d, ps = 64, 32768, 32, 8, 128, 256
    npages = b * sk // ps
    q = _gpu_randn((b, 1, h, d), 41)
    kp, vp = _gpu_randn((npages, ps, hk, d), 42), _gpu_randn((npages, ps, hk, d), 43)
    g = torch.Generator().manual_seed(44)
    table = torch.randperm(npages, generator=g).to(torch.int32).view(b, sk // ps).to(DEV)
    lens = (torch.randint(16384, sk + 1, (b,), generator=g, dtype=torch.int32) if ragged else torch.full((b,), sk, dtype=torch.int32)).to(DEV)
    outs = {}
    for ns in (0, 1, 4):
        outs[ns] = F.flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens, page_table=table, num_splits=ns, return_softmax_lse=True)[:2]
    for bi in (0, 17, 63):
        n = int(lens[bi])
        rows = table[bi].long()
        kd = kp[rows].reshape(1, sk, hk, d)[:, :n].cpu()
        vd = vp[rows].reshape(1, sk, hk, d)[:, :n].cpu()
        ref, lse_ref = A.attention_oracle(q[bi:bi + 1].cpu(), kd, vd)
        low, _ = A.attention_oracle(q[bi:bi + 1].cpu(), kd, vd, upcast=False, reorder_ops=True)
        for ns in outs:
            _check(outs[ns][0][bi:bi + 1], outs[ns][1][bi:bi + 1], ref, low, lse_ref)
    for ns in (1, 4):
        assert (outs[0][0].float() - outs[ns][0].float()).abs().max().item() <= 4e-3
        assert torch.allclose(outs[0][1], outs[ns][1], atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("num_splits", [0, 1, 4])
def test_decode_many_items_per_cta(num_splits):
    """Decode with more work items than SMs and 16 key blocks each, repeated: every CTA walks several items, the two
    issuing warps drift across item boundaries (regression test for a barrier-phase aliasing fault in the dual-stream
    decode kernel that only showed up with several long items per CTA)."""
    F = _iface()
    b, hk, sk, d = 40, 8, 2048, 128
    h = hk * 4
    g = torch.Generator(device=DEV).manual_seed(13)
    q = torch.randn(b, 1, h, d, generator=g, device=DEV, dtype=torch.bfloat16)
    k = torch.randn(b, sk, hk, d, generator=g, device=DEV, dtype=torch.bfloat16)
    v = torch.randn(b, sk, hk, d, generator=g, device=DEV, dtype=torch.bfloat16)
    lens = torch.randint(sk // 2, sk + 1, (b,), device=DEV, dtype=torch.int32, generator=g)
    first = None
    for _ in range(10):
        out, lse, *_ = F.flash_attn_with_kvcache(q, k, v, cache_seqlens=lens, num_splits=num_splits, return_softmax_lse=True)
        torch.cuda.synchronize()
        if first is None:
            first = (out.clone(), lse.clone())
        assert torch.equal(out, first[0]) and torch.equal(lse, first[1])
    for bi in (0, 17, 39):
        n = int(lens[bi])
        ref, lse_ref = A.attention_oracle(q[bi:bi + 1].cpu(), k[bi:bi + 1, :n].cpu(), v[bi:bi + 1, :n].cpu())
        low, _ = A.attention_oracle(q[bi:bi + 1].cpu(), k[bi:bi + 1, :n].cpu(), v[bi:bi + 1, :n].cpu(), upcast=False, reorder_ops=True)
        _check(out[bi:bi + 1], lse[bi:bi + 1], ref, low, lse_ref)


def test_cuda_graph_decode_with_scheduler_metadata():
    """hopper/test_flash_attn.py:879-929: get_scheduler_metadata + a CUDA-graph-captured decode call; the captured call
    must follow cache_seqlens updated in place between replays (no host sync, no allocation inside the C ABI)."""
    F = _iface()
    g = torch.Generator().manual_seed(12)
    b, h, hk, d, cap, ps = 4, 8, 2, 128, 2048, 256
    kd, vd, table, kp, vp = A.paged_kvcache(cap, ps, b, hk, d, d, torch.bfloat16, generator=g)
    q = torch.randn(b, 1, h, d, generator=g).to(torch.bfloat16)
    lens = torch.tensor([100, 2048, 777, 1], dtype=torch.int32)
    qd, kpd, vpd, td, ld = q.to(DEV), kp.to(DEV), vp.to(DEV), table.to(DEV), lens.to(DEV)
    meta = F.get_scheduler_metadata(b, 1, cap, h, hk, d, ld, qkv_dtype=torch.bfloat16, page_size=ps, num_splits=4)
    kw = dict(cache_seqlens=ld, page_table=td, scheduler_metadata=meta, num_splits=4, return_softmax_lse=True)
    F.flash_attn_with_kvcache(qd, kpd, vpd, **kw)  # warm-up outside the capture (function attributes, tensor maps)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out, lse, *_ = F.flash_attn_with_kvcache(qd, kpd, vpd, **kw)
    for new_lens in ([100, 2048, 777, 1], [2048, 5, 1300, 640]):
        ld.copy_(torch.tensor(new_lens, dtype=torch.int32))
        graph.replay()
        torch.cuda.synchronize()
        kl = torch.tensor(new_lens, dtype=torch.int32)
        ref, lse_ref = A.attention_oracle(q, kd, vd, k_lens=kl)
        low, _ = A.attention_oracle(q, kd, vd, k_lens=kl, upcast=False, reorder_ops=True)
        _check(out, lse, ref, low, lse_ref)


def test_unsupported_raises():
    F = _iface()
    q, k, v = (x.to(DEV) for x in _qkv(1, 64, 64, 2, 2, 512, torch.bfloat16))
    with pytest.raises(NotImplementedError):
        F.flash_attn_func(q, k, v)  # head dim > 256 (hopper/flash_api.cpp:710 has the same limit)
    q, k, v = _qkv(1, 64, 64, 2, 2, 64, torch.bfloat16)
    with pytest.raises((RuntimeError, NotImplementedError)):
        F.flash_attn_func(q, k, v)  # CPU tensors: no fallback


def test_prepare_scheduler_metadata():
    """fa100_prepare_scheduler (role of prepare_varlen_num_blocks, hopper/flash_prepare_scheduler.cu:43-215): batch
    elements ordered by key-block count, longest first; the prefix counts only row blocks that exist."""
    F = _iface()
    b, h, hk, d = 7, 8, 2, 128
    lens_q = torch.tensor([300, 0, 1024, 17, 512, 256, 257], dtype=torch.int32)
    lens_k = torch.tensor([900, 64, 1024, 4000, 0, 256, 257], dtype=torch.int32)
    cu_q = torch.nn.functional.pad(lens_q.cumsum(0), (1, 0)).to(torch.int32)
    cu_k = torch.nn.functional.pad(lens_k.cumsum(0), (1, 0)).to(torch.int32)
    meta = F.get_scheduler_metadata(b, int(lens_q.max()), int(lens_k.max()), h, hk, d, lens_k.to(DEV), qkv_dtype=torch.bfloat16,
                                    cu_seqlens_q=cu_q.to(DEV), causal=False, num_splits=1).cpu()
    assert meta[0].item() == 0xFA100
    body = meta[8:]
    rows_item = 512 if os.environ.get("FA100_PAIR") == "1" else 256              # query rows per work item (no pack-GQA here)
    m_blocks = [(int(x) + rows_item - 1) // rows_item for x in lens_q]
    # a tile past the end of the sequence has no key blocks: count the real 128-row tiles
    cost = [((int(q) + 127) // 128 if mb else 0) * ((int(k) + 127) // 128) for q, mb, k in zip(lens_q, m_blocks, lens_k)]
    order = sorted(range(b), key=lambda i: (-cost[i], i))
    items = [m_blocks[i] * h for i in order]
    assert body[1].item() == b and body[2:2 + b].tolist() == order
    pref = [0]
    for x in items:
        pref.append(pref[-1] + x)
    assert body[2 + b:2 + 2 * b + 1].tolist() == pref and body[0].item() == pref[-1]
    assert body[3 + 2 * b:3 + 3 * b].tolist() == [1] * b   # num_splits = 1: nothing to cut
    # and the forward that walks this list agrees with the oracle
    g = torch.Generator().manual_seed(77)
    q = torch.randn(int(cu_q[-1]), h, d, generator=g).to(torch.bfloat16)
    k = torch.randn(int(cu_k[-1]), hk, d, generator=g).to(torch.bfloat16)
    v = torch.randn(int(cu_k[-1]), hk, d, generator=g).to(torch.bfloat16)
    out, lse = F.flash_attn_varlen_func(q.to(DEV), k.to(DEV), v.to(DEV), cu_q.to(DEV), cu_k.to(DEV), int(lens_q.max()), int(lens_k.max()),
                                        return_attn_probs=True)
    for i in range(b):
        if lens_q[i] == 0:
            continue
        qs, ks = slice(int(cu_q[i]), int(cu_q[i + 1])), slice(int(cu_k[i]), int(cu_k[i + 1]))
        if lens_k[i] == 0:
            assert out[qs].abs().max().item() == 0
            continue
        ref, lse_ref = A.attention_oracle(q[qs][None], k[ks][None], v[ks][None])
        low, _ = A.attention_oracle(q[qs][None], k[ks][None], v[ks][None], upcast=False, reorder_ops=True)
        _check(out[qs][None], lse[:, qs][None], ref, low, lse_ref)


@pytest.mark.parametrize("page_size", [None, 256])
def test_ragged_decode_dynamic_splits(page_size):
    """Ragged decode with num_splits left to the library: the schedule kernel cuts the long sequences into several
    splits and leaves the short ones whole (hopper/flash_prepare_scheduler.cu:96-120); unused partials read as empty."""
    F = _iface()
    g = torch.Generator().manual_seed(91)
    b, h, hk, d, cap = 12, 8, 2, 128, 8192
    lens = torch.tensor([8192, 130, 64, 8000, 1, 512, 300, 4097, 128, 129, 2048, 7], dtype=torch.int32)
    q = torch.randn(b, 1, h, d, generator=g).to(torch.bfloat16)
    if page_size is None:
        kd = torch.randn(b, cap, hk, d, generator=g).to(torch.bfloat16)
        vd = torch.randn(b, cap, hk, d, generator=g).to(torch.bfloat16)
        kin, vin, table = kd, vd, None
    else:
        kd, vd, table, kin, vin = A.paged_kvcache(cap, page_size, b, hk, d, d, torch.bfloat16, generator=g)
    ref, lse_ref = A.attention_oracle(q, kd, vd, k_lens=lens)
    low, _ = A.attention_oracle(q, kd, vd, k_lens=lens, upcast=False, reorder_ops=True)
    out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kin.to(DEV), vin.to(DEV), cache_seqlens=lens.to(DEV),
                                             page_table=None if table is None else table.to(DEV), return_softmax_lse=True)
    _check(out, lse, ref, low, lse_ref)
    meta = F.get_scheduler_metadata(b, 1, cap, h, hk, d, lens.to(DEV), qkv_dtype=torch.bfloat16, page_size=page_size, num_splits=8).cpu()
    body = meta[8:]
    order, ns = body[2:2 + b].tolist(), body[3 + 2 * b:3 + 3 * b].tolist()
    assert order[0] == 0 and max(ns) > 1 and ns[order.index(4)] == 1 and all(1 <= x <= 4 for x in ns)   # 8 partials = 4 CTA-level splits
    out2, lse2, *_ = F.flash_attn_with_kvcache(q.to(DEV), kin.to(DEV), vin.to(DEV), cache_seqlens=lens.to(DEV),
                                               page_table=None if table is None else table.to(DEV), scheduler_metadata=meta.to(DEV),
                                               num_splits=8, return_softmax_lse=True)
    _check(out2, lse2, ref, low, lse_ref)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("sq,sk", [(256, 512), (1024, 1024), (700, 500), (1500, 3000)])
def test_two_cta_variant(sq, sk, causal, monkeypatch):
    """The cta_group::2 kernel (Fa100FwdParams.variant = 1 via FA100_PAIR=1): CTA pairs, M = 256 MMAs, each CTA holding half
    of every K / V tile.  Same oracle, same tolerance; also bitwise equal to the 1-CTA kernel (same per-row arithmetic)."""
    F = _iface()
    q, k, v = _qkv(2, sq, sk, 8, 2, 128, torch.bfloat16, seed=5)
    ref, lse_ref = A.attention_oracle(q, k, v, causal=causal)
    low, _ = A.attention_oracle(q, k, v, causal=causal, upcast=False, reorder_ops=True)
    monkeypatch.setenv("FA100_PAIR", "0")
    base, _ = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), causal=causal, return_attn_probs=True)
    monkeypatch.setenv("FA100_PAIR", "1")
    for num_splits in (1, 3):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), causal=causal, num_splits=num_splits, return_attn_probs=True)
        _check(out, lse, ref, low, lse_ref)
        if num_splits == 1:
            assert torch.equal(out, base)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float8_e4m3fn])
@pytest.mark.parametrize("kw", [dict(), dict(causal=True), dict(window_size=(100, 20)), dict(softcap=15.0, causal=True)])
@pytest.mark.parametrize("sq,sk,d", [(256, 512, 128), (1024, 1024, 128), (700, 500, 64), (1500, 3000, 128), (113, 203, 96)])
def test_four_role_variant(sq, sk, d, kw, dtype, monkeypatch):
    """The four-role prefill mainloop of csrc/fa100_fwd2.cuh (FA100_FOUR_ROLE=1 -> Fa100FwdParams.variant bit 32): one MMA
    warp issuing PV_t, QK_t in order with P_t overlaid on S_t, separate softmax and correction warpgroups (the role split
    north_star names), P published in two pieces.  Same oracle and tolerance as the default mainloop, every epilogue form
    (TMA store, pack-GQA vector stores, fp32 partials + combine)."""
    F = _iface()
    fp8 = dtype == torch.float8_e4m3fn
    if fp8 and d == 96:
        d = 64
    g = torch.Generator().manual_seed(9)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16).to(dtype)  # noqa: E731
    q, k, v = mk(2, sq, 8, d), mk(2, sk, 2, d), mk(2, sk, 2, d)
    qr, kr, vr = (x.to(torch.bfloat16) if fp8 else x for x in (q, k, v))
    ref, lse_ref = A.attention_oracle(qr, kr, vr, **kw)
    low, _ = A.attention_oracle(qr, kr, vr, upcast=False, reorder_ops=True, intermediate_dtype=dtype if fp8 else None, **kw)
    monkeypatch.setenv("FA100_FOUR_ROLE", "1")
    for pack_gqa, num_splits in ((False, 1), (True, 1), (False, 3)):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), pack_gqa=pack_gqa, num_splits=num_splits, return_attn_probs=True, **kw)
        _check(out, lse, ref, low, lse_ref, rtol=4.0 if fp8 else (3.0 if kw.get("softcap") else 2.0))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float8_e4m3fn])
@pytest.mark.parametrize("kw", [dict(), dict(causal=True), dict(window_size=(100, 20)), dict(softcap=15.0, causal=True)])
@pytest.mark.parametrize("sq,sk,d", [(256, 512, 128), (1024, 1024, 128), (700, 500, 64), (1500, 3000, 128), (113, 203, 96), (300, 700, 256)])
def test_split_rows_variant(sq, sk, d, kw, dtype, monkeypatch):
    """The split-row prefill kernel of csrc/fa100_fwd_split.cuh (FA100_SPLIT_ROWS=1 -> Fa100FwdParams.variant bit 64): two
    softmax threads per query row, each on 64 of a block's 128 key columns and half of the O columns, identical running
    maxima without a hand-shake, row sums joined once per work item.  Same oracle and tolerance as the default kernel,
    every epilogue form (TMA store, pack-GQA vector stores, fp32 partials + combine)."""
    F = _iface()
    fp8 = dtype == torch.float8_e4m3fn
    if fp8 and d == 96:
        d = 64
    # d = 256: the one-tile (head dim > 128) geometry, where the eight softmax warps share the single Q tile
    g = torch.Generator().manual_seed(10)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16).to(dtype)  # noqa: E731
    q, k, v = mk(2, sq, 8, d), mk(2, sk, 2, d), mk(2, sk, 2, d)
    qr, kr, vr = (x.to(torch.bfloat16) if fp8 else x for x in (q, k, v))
    ref, lse_ref = A.attention_oracle(qr, kr, vr, **kw)
    low, _ = A.attention_oracle(qr, kr, vr, upcast=False, reorder_ops=True, intermediate_dtype=dtype if fp8 else None, **kw)
    monkeypatch.setenv("FA100_SPLIT_ROWS", "1")
    for pack_gqa, num_splits in ((False, 1), (True, 1), (False, 3)):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), pack_gqa=pack_gqa, num_splits=num_splits, return_attn_probs=True, **kw)
        _check(out, lse, ref, low, lse_ref, rtol=4.0 if fp8 else (3.0 if kw.get("softcap") else 2.0))
        again, _ = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), pack_gqa=pack_gqa, num_splits=num_splits, return_attn_probs=True, **kw)
        assert torch.equal(out, again)  # the halves of a row agree on every rescale decision: bitwise repeatable
