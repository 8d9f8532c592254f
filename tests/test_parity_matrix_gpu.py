"""Parity-matrix cells the reference's tests cover and tests/test_fwd_gpu.py did not (round-1 review): FP8 through every
feature path (hopper/test_flash_attn.py:63,290,602 run e4m3 through local windows, softcap, varlen, the KV cache incl.
paged / leftpad / batch_idx / splits / append + rotary), kvcache x cu_seqlens_q (:626), leftpad x append (:622 x :607), the
remaining shapes of the dense matrix (:78-99) with head dims 96 / 192 / 256, 1000 bitwise repeats under a large allocation
(:1043-1102), garbage past cache_seqlens, two devices from one process, and the schedule-header contract.

All through the C ABI (flash_attn_interface -> torch.ops.flash_attn_3 -> libfa100.so) against the pinned CPU oracle.
Tolerances: bf16/fp16 2x the low-precision oracle's error (+ atol), FP8 4x (hopper/test_flash_attn.py:986-989), LSE 1e-3."""
import math
import os

import pytest
import torch

import attention_oracle as A

pytestmark = pytest.mark.gpu
DEV = "cuda"
E4M3 = torch.float8_e4m3fn


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


def _fp8(*shape, g):
    """e4m3 test data the way the reference draws it: randn in bf16, rounded to e4m3 (hopper/test_flash_attn.py:131-149);
    returns (e4m3 tensor, the same values in bf16 for the oracle)."""
    x = torch.randn(*shape, generator=g).to(torch.bfloat16).to(E4M3)
    return x, x.to(torch.bfloat16)


def _descales(b, hk, g):
    return [torch.rand(b, hk, generator=g) * 2 for _ in range(3)]


def _oracle_fp8(q, k, v, ds, **kw):
    okw = dict(kw, q_descale=ds[0], k_descale=ds[1], v_descale=ds[2])
    ref, lse_ref = A.attention_oracle(q, k, v, **okw)
    low, _ = A.attention_oracle(q, k, v, upcast=False, reorder_ops=True, intermediate_dtype=E4M3, **okw)
    return ref, low, lse_ref


# ----------------------------------------------------------------------------- FP8 x dense features
@pytest.mark.parametrize("kw", [dict(window_size=(100, 50)), dict(window_size=(300, 0)), dict(window_size=(-1, 30)),
                                dict(attention_chunk=128), dict(softcap=15.0), dict(softcap=15.0, causal=True),
                                dict(causal=True), dict()])
@pytest.mark.parametrize("mha_type", ["mha", "gqa", "mqa"])
@pytest.mark.parametrize("sq,sk", [(500, 700), (700, 500), (64, 1500), (1, 333), (1024, 1024)])
def test_fp8_masks_softcap_gqa_splits(sq, sk, mha_type, kw):
    """e4m3 x {local window, chunk, softcap, causal} x {mha, gqa, mqa} x {pack_gqa} x {split-KV + combine}."""
    F = _iface()
    g = torch.Generator().manual_seed(21)
    b, h, d = 2, 4, 128
    hk = {"mha": 4, "gqa": 2, "mqa": 1}[mha_type]
    (q8, q), (k8, k), (v8, v) = _fp8(b, sq, h, d, g=g), _fp8(b, sk, hk, d, g=g), _fp8(b, sk, hk, d, g=g)
    ds = _descales(b, hk, g)
    ref, low, lse_ref = _oracle_fp8(q, k, v, ds, **kw)
    for pack_gqa, num_splits in ((False, 1), (True, 1), (None, 3)):
        out, lse = F.flash_attn_func(q8.to(DEV), k8.to(DEV), v8.to(DEV), q_descale=ds[0].to(DEV), k_descale=ds[1].to(DEV),
                                     v_descale=ds[2].to(DEV), pack_gqa=pack_gqa, num_splits=num_splits, return_attn_probs=True, **kw)
        _check(out, lse, ref, low, lse_ref, rtol=4.0)


# ----------------------------------------------------------------------------- FP8 x varlen
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("lens_q,lens_k", [([100, 257, 31], [100, 300, 77]), ([128, 5, 300, 1], [128, 0, 511, 1000]),
                                           ([1, 1, 1, 1, 1], [7, 130, 250, 1, 600])])
def test_fp8_varlen(lens_q, lens_k, causal):
    F = _iface()
    g = torch.Generator().manual_seed(22)
    h, hk, d = 4, 2, 128
    b = len(lens_q)
    cq = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), dtype=torch.int32)
    ck = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), dtype=torch.int32)
    (q8, q), (k8, k), (v8, v) = _fp8(int(cq[-1]), h, d, g=g), _fp8(int(ck[-1]), hk, d, g=g), _fp8(int(ck[-1]), hk, d, g=g)
    ds = _descales(b, hk, g)
    refs, lows, lses = [], [], []
    for i in range(b):
        qi, ki, vi = q[cq[i]:cq[i + 1]][None], k[ck[i]:ck[i + 1]][None], v[ck[i]:ck[i + 1]][None]
        if lens_k[i] == 0:
            refs.append(torch.zeros_like(qi[0])); lows.append(torch.zeros_like(qi[0]))
            lses.append(torch.full((h, lens_q[i]), float("-inf")))
            continue
        r, lo, l = _oracle_fp8(qi, ki, vi, [x[i:i + 1] for x in ds], causal=causal)
        refs.append(r[0]); lows.append(lo[0]); lses.append(l[0])
    for pack_gqa, num_splits in ((False, 1), (True, 2)):
        out, lse = F.flash_attn_varlen_func(q8.to(DEV), k8.to(DEV), v8.to(DEV), cq.to(DEV), ck.to(DEV), max(lens_q), max(lens_k),
                                            q_descale=ds[0].to(DEV), k_descale=ds[1].to(DEV), v_descale=ds[2].to(DEV),
                                            causal=causal, pack_gqa=pack_gqa, num_splits=num_splits, return_attn_probs=True)
        _check(out, lse, torch.cat(refs), torch.cat(lows), torch.cat(lses, dim=1), rtol=4.0)


# ----------------------------------------------------------------------------- FP8 x KV cache (dense / paged / batch_idx / leftpad / splits)
@pytest.mark.parametrize("page_size", [None, 1, 4, 128, 256])
@pytest.mark.parametrize("sq,sk", [(1, 128), (1, 339), (3, 1024), (64, 800), (16, 3000)])
@pytest.mark.parametrize("causal,local", [(False, False), (True, False), (False, True)])
def test_fp8_kvcache(sq, sk, page_size, causal, local):
    F = _iface()
    g = torch.Generator().manual_seed(23)
    b, h, hk, d = 3, 6, 3, 128
    q8, q = _fp8(b, sq, h, d, g=g)
    lens = torch.randint(max(1, sk // 2), sk + 1, (b,), generator=g, dtype=torch.int32).clamp_min(sq)
    if page_size is None:
        (k8, kd), (v8, vd) = _fp8(b, sk, hk, d, g=g), _fp8(b, sk, hk, d, g=g)
        kin, vin, table = k8, v8, None
    else:
        skp = (sk + page_size - 1) // page_size * page_size
        kd, vd, table, kin, vin = A.paged_kvcache(skp, page_size, b, hk, d, d, torch.bfloat16, generator=g)
        kin, vin = kin.to(E4M3), vin.to(E4M3)
        kd, vd = kd.to(E4M3).to(torch.bfloat16), vd.to(E4M3).to(torch.bfloat16)
    ds = _descales(b, hk, g)
    kw = dict(causal=causal, window_size=(37, 11) if local else (-1, -1))
    ref, low, lse_ref = _oracle_fp8(q, kd, vd, ds, k_lens=lens, **kw)
    for num_splits in (0, 1, 3):
        out, lse, *_ = F.flash_attn_with_kvcache(q8.to(DEV), kin.to(DEV), vin.to(DEV), cache_seqlens=lens.to(DEV),
                                                 page_table=None if table is None else table.to(DEV), num_splits=num_splits,
                                                 q_descale=ds[0].to(DEV), k_descale=ds[1].to(DEV), v_descale=ds[2].to(DEV),
                                                 return_softmax_lse=True, **kw)
        _check(out, lse, ref, low, lse_ref, rtol=4.0)


def test_fp8_kvcache_batch_idx_and_leftpad():
    F = _iface()
    g = torch.Generator().manual_seed(24)
    b, bc, sq, sk, h, hk, d = 3, 7, 2, 600, 4, 2, 128
    q8, q = _fp8(b, sq, h, d, g=g)
    (k8, kc), (v8, vc) = _fp8(bc, sk, hk, d, g=g), _fp8(bc, sk, hk, d, g=g)
    idx = torch.tensor([5, 0, 3], dtype=torch.int32)
    lens = torch.tensor([600, 257, 400], dtype=torch.int32)
    pad = torch.tensor([0, 100, 37], dtype=torch.int32)
    ds = _descales(b, hk, g)
    out, lse, *_ = F.flash_attn_with_kvcache(q8.to(DEV), k8.to(DEV), v8.to(DEV), cache_seqlens=lens.to(DEV), cache_batch_idx=idx.to(DEV),
                                             cache_leftpad=pad.to(DEV), causal=True, q_descale=ds[0].to(DEV), k_descale=ds[1].to(DEV),
                                             v_descale=ds[2].to(DEV), return_softmax_lse=True)
    ref, low, lse_ref = _oracle_fp8(q, kc[idx.long()], vc[idx.long()], ds, k_lens=lens - pad, key_leftpad=pad, causal=True)
    _check(out, lse, ref, low, lse_ref, rtol=4.0)


# ----------------------------------------------------------------------------- FP8 x append-KV x rotary
@pytest.mark.parametrize("rotary", [None, "interleaved", "halves", "partial"])
@pytest.mark.parametrize("layout", ["dense", "paged64", "varlen_new"])
@pytest.mark.parametrize("mask", ["none", "causal"])
@pytest.mark.parametrize("sq,s_new", [(1, 1), (5, 5), (3, 130)])
def test_fp8_kvcache_append_rotary(sq, s_new, mask, layout, rotary):
    """e4m3 q / cache / k_new / v_new with cos / sin in e4m3 as well (the reference requires rotary_cos.dtype == q.dtype,
    hopper/flash_api.cpp:1052-1070; its test casts cos / sin to e4m3, hopper/test_flash_attn.py:616-618,877-878).  Cache:
    V bit-exact, K bit-exact without rotary and within the reference's 1e-1 with it (:972-984)."""
    F = _iface()
    g = torch.Generator().manual_seed(31)
    b, h, hk, d, cap = 3, 4, 2, 128, 512
    q8, q = _fp8(b, sq, h, d, g=g)
    kn8, kn = _fp8(b, s_new, hk, d, g=g)
    vn8, vn = _fp8(b, s_new, hk, d, g=g)
    lens = torch.randint(0, cap - s_new + 1, (b,), generator=g, dtype=torch.int32)
    new_lens = torch.full((b,), s_new, dtype=torch.int32)
    if layout == "varlen_new":
        new_lens = torch.randint(1, s_new + 1, (b,), generator=g, dtype=torch.int32)
        new_lens[0] = s_new
    causal = mask == "causal"
    if causal:
        lens = lens.clamp_min(sq)
    table = None
    if layout.startswith("paged"):
        ps = int(layout[5:])
        kd, vd, table, kc, vc = A.paged_kvcache(cap, ps, b, hk, d, d, torch.bfloat16, generator=g)
        kc, vc = kc.to(E4M3), vc.to(E4M3)
        kd, vd = kd.to(E4M3).to(torch.bfloat16).clone(), vd.to(E4M3).to(torch.bfloat16).clone()
    else:
        (kc, kd), (vc, vd) = _fp8(b, cap, hk, d, g=g), _fp8(b, cap, hk, d, g=g)
        kd, vd = kd.clone(), vd.clone()
    cos = sin = None
    q_ref, kn_ref = q, kn
    if rotary is not None:
        rd = 64 if rotary == "partial" else d
        ang = torch.rand(cap + 64, rd // 2, generator=g) * 2 * math.pi
        cos8, sin8 = torch.cos(ang).to(torch.bfloat16).to(E4M3), torch.sin(ang).to(torch.bfloat16).to(E4M3)
        cos, sin = cos8.to(torch.bfloat16), sin8.to(torch.bfloat16)
        inter = rotary == "interleaved"
        # the kernel rounds the rotated rows to e4m3 once (fp32 math); the oracle's bf16 result is rounded the same way
        kn_ref = A.rotary_oracle(kn.float(), cos, sin, lens, interleaved=inter).to(E4M3).to(torch.bfloat16)
        if causal:
            q_ref = A.rotary_oracle(q.float(), cos, sin, lens, interleaved=inter).to(E4M3).to(torch.bfloat16)
        else:
            q_ref = A.rotary_oracle(q.float().reshape(b, 1, sq * h, d), cos, sin, lens, interleaved=inter).reshape(b, sq, h, d).to(E4M3).to(torch.bfloat16)
    for i in range(b):
        n = int(new_lens[i])
        kd[i, int(lens[i]):int(lens[i]) + n] = kn_ref[i, :n]
        vd[i, int(lens[i]):int(lens[i]) + n] = vn[i, :n]
    tot = lens + new_lens
    ds = _descales(b, hk, g)
    ref, low, lse_ref = _oracle_fp8(q_ref, kd, vd, ds, k_lens=tot, causal=causal)
    kc_dev, vc_dev = kc.to(DEV), vc.to(DEV)
    if layout == "varlen_new":
        cu_new = torch.nn.functional.pad(new_lens.cumsum(0), (1, 0)).to(torch.int32)
        kn_in = torch.cat([kn8[i, :int(new_lens[i])] for i in range(b)]).to(DEV)
        vn_in = torch.cat([vn8[i, :int(new_lens[i])] for i in range(b)]).to(DEV)
        extra = dict(cu_seqlens_k_new=cu_new.to(DEV))
    else:
        kn_in, vn_in, extra = kn8.to(DEV), vn8.to(DEV), {}
    out, lse, *_ = F.flash_attn_with_kvcache(
        q8.to(DEV), kc_dev, vc_dev, k=kn_in, v=vn_in, rotary_cos=None if cos is None else cos8.to(DEV),
        rotary_sin=None if sin is None else sin8.to(DEV), cache_seqlens=lens.to(DEV), page_table=None if table is None else table.to(DEV),
        rotary_interleaved=(rotary == "interleaved"), causal=causal, q_descale=ds[0].to(DEV), k_descale=ds[1].to(DEV),
        v_descale=ds[2].to(DEV), return_softmax_lse=True, **extra)
    # the rotated q / k_new may differ from the oracle's by one e4m3 rounding step on a few elements: the reference allows
    # 1e-1 on the cache (hopper/test_flash_attn.py:982-984); the output keeps the FP8 multiplier
    _check(out, lse, ref, low, lse_ref, rtol=4.0 if rotary is None else 6.0)
    if table is not None:
        kd_dev = kc_dev.cpu().to(torch.bfloat16)[table.flatten().long()].view(b, -1, hk, d)[:, :cap]
        vd_dev = vc_dev.cpu().to(torch.bfloat16)[table.flatten().long()].view(b, -1, hk, d)[:, :cap]
    else:
        kd_dev, vd_dev = kc_dev.cpu().to(torch.bfloat16), vc_dev.cpu().to(torch.bfloat16)
    assert torch.equal(vd_dev, vd)
    if rotary is None:
        assert torch.equal(kd_dev, kd)
    else:
        assert torch.allclose(kd_dev.float(), kd.float(), rtol=1e-1, atol=1e-1)


# ----------------------------------------------------------------------------- kvcache x cu_seqlens_q, leftpad x append
@pytest.mark.parametrize("dtype", [torch.bfloat16, E4M3])
@pytest.mark.parametrize("page_size", [None, 64])
@pytest.mark.parametrize("causal", [False, True])
def test_kvcache_varlen_q(causal, page_size, dtype):
    """flash_attn_with_kvcache with a packed (cu_seqlens_q) query against a batched cache (hopper/test_flash_attn.py:626)."""
    F = _iface()
    g = torch.Generator().manual_seed(41)
    b, h, hk, d, sk = 4, 6, 2, 128, 900
    lens_q = [7, 64, 1, 33]
    cq = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), dtype=torch.int32)
    fp8 = dtype == E4M3
    mk = (lambda *s: _fp8(*s, g=g)) if fp8 else (lambda *s: (lambda x: (x, x))(torch.randn(*s, generator=g).to(dtype)))
    qin, q = mk(int(cq[-1]), h, d)
    lens = torch.randint(sk // 2, sk + 1, (b,), generator=g, dtype=torch.int32)
    if page_size is None:
        (kin, kd), (vin, vd) = mk(b, sk, hk, d), mk(b, sk, hk, d)
        table = None
    else:
        skp = (sk + page_size - 1) // page_size * page_size
        kd, vd, table, kin, vin = A.paged_kvcache(skp, page_size, b, hk, d, d, torch.bfloat16, generator=g)
        if fp8:
            kin, vin = kin.to(E4M3), vin.to(E4M3)
            kd, vd = kd.to(E4M3).to(torch.bfloat16), vd.to(E4M3).to(torch.bfloat16)
    out, lse, *_ = F.flash_attn_with_kvcache(qin.to(DEV), kin.to(DEV), vin.to(DEV), cache_seqlens=lens.to(DEV),
                                             page_table=None if table is None else table.to(DEV), cu_seqlens_q=cq.to(DEV),
                                             max_seqlen_q=max(lens_q), causal=causal, return_softmax_lse=True)
    for i in range(b):
        qs = slice(int(cq[i]), int(cq[i + 1]))
        kw = dict(causal=causal, k_lens=lens[i:i + 1])
        ref, lse_ref = A.attention_oracle(q[qs][None], kd[i:i + 1], vd[i:i + 1], **kw)
        low, _ = A.attention_oracle(q[qs][None], kd[i:i + 1], vd[i:i + 1], upcast=False, reorder_ops=True,
                                    intermediate_dtype=E4M3 if fp8 else None, **kw)
        _check(out[qs][None], lse[:, qs][None], ref, low, lse_ref, rtol=4.0 if fp8 else 2.0)


@pytest.mark.parametrize("rotary", [False, True])
@pytest.mark.parametrize("causal", [False, True])
def test_kvcache_leftpad_append(causal, rotary):
    """cache_leftpad together with appended keys (hopper/test_flash_attn.py:622 x :607; hopper/seqlen.h:82-92: the new rows
    land at cache_seqlens, which counts from the start of the buffer, and rotary positions do not see the left pad)."""
    F = _iface()
    g = torch.Generator().manual_seed(42)
    b, sq, s_new, h, hk, d, cap = 3, 4, 4, 4, 2, 128, 400
    dt = torch.bfloat16
    q = torch.randn(b, sq, h, d, generator=g).to(dt)
    kn, vn = torch.randn(b, s_new, hk, d, generator=g).to(dt), torch.randn(b, s_new, hk, d, generator=g).to(dt)
    kc, vc = torch.randn(b, cap, hk, d, generator=g).to(dt), torch.randn(b, cap, hk, d, generator=g).to(dt)
    lens = torch.tensor([200, 57, 396], dtype=torch.int32)
    pad = torch.tensor([0, 30, 120], dtype=torch.int32)
    cos = sin = None
    q_ref, kn_ref = q, kn
    if rotary:
        ang = torch.rand(cap + 8, d // 2, generator=g) * 2 * math.pi
        cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)
        kn_ref = A.rotary_oracle(kn, cos, sin, lens, interleaved=False)
        q_ref = A.rotary_oracle(q, cos, sin, lens, interleaved=False) if causal else \
            A.rotary_oracle(q.reshape(b, 1, sq * h, d), cos, sin, lens, interleaved=False).reshape(b, sq, h, d)
    kd, vd = kc.clone(), vc.clone()
    for i in range(b):
        kd[i, int(lens[i]):int(lens[i]) + s_new] = kn_ref[i]
        vd[i, int(lens[i]):int(lens[i]) + s_new] = vn[i]
    tot = lens + s_new
    kw = dict(causal=causal, k_lens=tot - pad, key_leftpad=pad)
    ref, lse_ref = A.attention_oracle(q_ref, kd, vd, **kw)
    low, _ = A.attention_oracle(q_ref, kd, vd, upcast=False, reorder_ops=True, **kw)
    kc_dev, vc_dev = kc.to(DEV), vc.to(DEV)
    out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kc_dev, vc_dev, k=kn.to(DEV), v=vn.to(DEV), cache_seqlens=lens.to(DEV),
                                             cache_leftpad=pad.to(DEV), rotary_cos=None if cos is None else cos.to(DEV),
                                             rotary_sin=None if sin is None else sin.to(DEV), rotary_interleaved=False,
                                             causal=causal, return_softmax_lse=True)
    _check(out, lse, ref, low, lse_ref)
    assert torch.equal(vc_dev.cpu(), vd)
    assert torch.allclose(kc_dev.cpu().float(), kd.float(), rtol=1e-3, atol=1e-3)


def test_rotary_window_larger_than_sequence():
    """A window that covers the whole problem is no window (hopper/flash_api.cpp:796-805): the query rows then all rotate at
    cache_seqlens (FixedPosition), not at cache_seqlens + row."""
    F = _iface()
    g = torch.Generator().manual_seed(43)
    b, sq, h, hk, d, cap = 2, 4, 4, 2, 128, 256
    dt = torch.bfloat16
    q = torch.randn(b, sq, h, d, generator=g).to(dt)
    kn, vn = torch.randn(b, sq, hk, d, generator=g).to(dt), torch.randn(b, sq, hk, d, generator=g).to(dt)
    kc, vc = torch.randn(b, cap, hk, d, generator=g).to(dt), torch.randn(b, cap, hk, d, generator=g).to(dt)
    lens = torch.tensor([100, 31], dtype=torch.int32)
    ang = torch.rand(cap + 8, d // 2, generator=g) * 2 * math.pi
    cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)
    kn_ref = A.rotary_oracle(kn, cos, sin, lens)
    q_ref = A.rotary_oracle(q.reshape(b, 1, sq * h, d), cos, sin, lens).reshape(b, sq, h, d)  # fixed position
    kd, vd = kc.clone(), vc.clone()
    for i in range(b):
        kd[i, int(lens[i]):int(lens[i]) + sq] = kn_ref[i]
        vd[i, int(lens[i]):int(lens[i]) + sq] = vn[i]
    ref, lse_ref = A.attention_oracle(q_ref, kd, vd, k_lens=lens + sq)
    low, _ = A.attention_oracle(q_ref, kd, vd, k_lens=lens + sq, upcast=False, reorder_ops=True)
    out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kc.to(DEV), vc.to(DEV), k=kn.to(DEV), v=vn.to(DEV), cache_seqlens=lens.to(DEV),
                                             rotary_cos=cos.to(DEV), rotary_sin=sin.to(DEV), rotary_interleaved=False,
                                             window_size=(cap, cap), return_softmax_lse=True)
    _check(out, lse, ref, low, lse_ref)


# ----------------------------------------------------------------------------- the rest of the dense matrix
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("mha_type", ["mha", "mqa", "gqa"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 96, 128, 192, 256])
@pytest.mark.parametrize("sq,sk", [(128, 192), (113, 128), (1024, 1023)])
def test_output_remaining_shapes(sq, sk, d, causal, mha_type, dtype):
    """The shapes of hopper/test_flash_attn.py:78-99 that test_output leaves out, over d in {64, 96, 128, 192, 256} x
    mha / mqa / gqa x pack_gqa x num_splits (the large ones are in test_output_large_shapes)."""
    F = _iface()
    h = 6
    hk = {"mha": 6, "mqa": 1, "gqa": 2}[mha_type]
    g = torch.Generator().manual_seed(51)
    q, k, v = (torch.randn(3, s_, h_, d, generator=g).to(dtype) for s_, h_ in ((sq, h), (sk, hk), (sk, hk)))
    ref, lse_ref = A.attention_oracle(q, k, v, causal=causal)
    low, _ = A.attention_oracle(q, k, v, causal=causal, upcast=False, reorder_ops=True)
    for pack_gqa, num_splits in ((False, 1), (True, 1), (False, 3), (True, 3)):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), causal=causal, pack_gqa=pack_gqa, num_splits=num_splits,
                                     return_attn_probs=True)
        _check(out, lse, ref, low, lse_ref)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d,mha_type", [(128, "gqa"), (64, "mha"), (256, "mqa"), (96, "gqa"), (192, "mha")])
@pytest.mark.parametrize("sq,sk", [(4096, 4096), (4224, 4224)])
def test_output_large_shapes(sq, sk, d, mha_type, causal):
    """(4096, 4096) and (4224, 4224) of hopper/test_flash_attn.py:97-98: enough tiles for every CTA's persistent loop to run
    several work items.  The CPU oracle runs on one batch element."""
    F = _iface()
    h = 6
    hk = {"mha": 6, "mqa": 1, "gqa": 2}[mha_type]
    g = torch.Generator(device=DEV).manual_seed(52)
    q, k, v = (torch.randn(2, s_, h_, d, generator=g, device=DEV, dtype=torch.bfloat16) for s_, h_ in ((sq, h), (sk, hk), (sk, hk)))
    qc, kc, vc = q[1:].cpu(), k[1:].cpu(), v[1:].cpu()
    ref, lse_ref = A.attention_oracle(qc, kc, vc, causal=causal)
    low, _ = A.attention_oracle(qc, kc, vc, causal=causal, upcast=False, reorder_ops=True)
    for pack_gqa, num_splits in ((False, 1), (True, 3)):
        out, lse = F.flash_attn_func(q, k, v, causal=causal, pack_gqa=pack_gqa, num_splits=num_splits, return_attn_probs=True)
        _check(out[1:], lse[1:], ref, low, lse_ref)


# ----------------------------------------------------------------------------- repeatability, robustness, boundary
def test_bitwise_repeatable_1000_under_large_allocation():
    """hopper/test_flash_attn.py:1043-1102: 1000 calls of the same problem must be bitwise identical, with most of the HBM
    held by a dummy allocation (the reference takes 70 GB of its 80; here 120 GB of 180) so that a stray write or a
    dependence on uninitialised memory shows."""
    F = _iface()
    free, _ = torch.cuda.mem_get_info()
    hold_gb = min(120, int(free / 2 ** 30) - 16)
    dummy = torch.empty(max(hold_gb, 1) * 2 ** 30, dtype=torch.uint8, device=DEV) if hold_gb > 0 else None
    g = torch.Generator(device=DEV).manual_seed(61)
    b, s, h, hk, d = 2, 2048, 8, 2, 128
    q, k, v = (torch.randn(b, s, h_, d, generator=g, device=DEV, dtype=torch.bfloat16) for h_ in (h, hk, hk))
    out0, lse0 = F.flash_attn_func(q, k, v, causal=True, return_attn_probs=True)
    out0, lse0 = out0.clone(), lse0.clone()
    bad = torch.zeros((), dtype=torch.int64, device=DEV)
    for _ in range(1000):
        out, lse = F.flash_attn_func(q, k, v, causal=True, return_attn_probs=True)
        bad += (out.view(torch.int16) != out0.view(torch.int16)).sum() + (lse.view(torch.int32) != lse0.view(torch.int32)).sum()
    assert bad.item() == 0
    # and the split / decode path (atomic-free partials + combine)
    qd = q[:, :1].contiguous()
    lens = torch.tensor([2048, 1500], dtype=torch.int32, device=DEV)
    o0 = F.flash_attn_with_kvcache(qd, k, v, cache_seqlens=lens).clone()
    bad.zero_()
    for _ in range(1000):
        bad += (F.flash_attn_with_kvcache(qd, k, v, cache_seqlens=lens).view(torch.int16) != o0.view(torch.int16)).sum()
    assert bad.item() == 0
    del dummy


@pytest.mark.parametrize("poison", [float("nan"), float("inf")])
@pytest.mark.parametrize("page_size", [None, 64])
@pytest.mark.parametrize("sq", [1, 5, 200])
def test_garbage_past_cache_seqlens(sq, page_size, poison):
    """Key / value rows past cache_seqlens are real memory, not TMA zero fill -- an uninitialised serving cache can hold NaN
    or Inf there.  Their probabilities are 0, and the kernel zeroes those V rows in shared memory so that 0 x NaN cannot
    poison the tile (the K side is handled by the mask, which replaces the score whatever it is)."""
    F = _iface()
    g = torch.Generator().manual_seed(62)
    b, h, hk, d, sk = 3, 8, 2, 128, 640
    q = torch.randn(b, sq, h, d, generator=g).to(torch.bfloat16)
    lens = torch.tensor([333, 640, 129], dtype=torch.int32).clamp_min(sq)
    if page_size is None:
        kd = torch.randn(b, sk, hk, d, generator=g).to(torch.bfloat16)
        vd = torch.randn(b, sk, hk, d, generator=g).to(torch.bfloat16)
        kin, vin, table = kd.clone(), vd.clone(), None
        for i in range(b):
            kin[i, int(lens[i]):] = poison
            vin[i, int(lens[i]):] = poison
    else:
        kd, vd, table, kin, vin = A.paged_kvcache(sk, page_size, b, hk, d, d, torch.bfloat16, generator=g)
        kin, vin = kin.clone(), vin.clone()
        used = torch.zeros(kin.shape[0], page_size, dtype=torch.bool)
        for i in range(b):
            for r in range(int(lens[i])):
                used[int(table[i, r // page_size]), r % page_size] = True
        kin[~used] = poison
        vin[~used] = poison
    for causal in (False, True):
        ref, lse_ref = A.attention_oracle(q, kd, vd, k_lens=lens, causal=causal)
        low, _ = A.attention_oracle(q, kd, vd, k_lens=lens, causal=causal, upcast=False, reorder_ops=True)
        for num_splits in (0, 1):
            out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kin.to(DEV), vin.to(DEV), cache_seqlens=lens.to(DEV),
                                                     page_table=None if table is None else table.to(DEV), causal=causal,
                                                     num_splits=num_splits, return_softmax_lse=True)
            _check(out, lse, ref, low, lse_ref)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (run with gpurun --gpus 2)")
def test_two_devices_one_process():
    """The one-time cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device: a second GPU used from the
    same process must get its own opt-in (225 KB of dynamic shared memory), for the prefill, decode and hdim-256 kernels."""
    F = _iface()
    g = torch.Generator().manual_seed(63)
    q, k, v = (torch.randn(2, 300, h_, 128, generator=g).to(torch.bfloat16) for h_ in (8, 2, 2))
    ref, lse_ref = A.attention_oracle(q, k, v, causal=True)
    low, _ = A.attention_oracle(q, k, v, causal=True, upcast=False, reorder_ops=True)
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        out, lse = F.flash_attn_func(q.to(dev), k.to(dev), v.to(dev), causal=True, return_attn_probs=True)
        torch.cuda.synchronize(dev)
        assert out.device == torch.device(dev)
        _check(out, lse, ref, low, lse_ref)
        lens = torch.tensor([300, 77], dtype=torch.int32, device=dev)
        o2 = F.flash_attn_with_kvcache(q[:, :1].to(dev), k.to(dev), v.to(dev), cache_seqlens=lens)
        torch.cuda.synchronize(dev)
        r2, _ = A.attention_oracle(q[:, :1], k, v, k_lens=lens.cpu())
        assert (o2.float().cpu() - r2.float()).abs().max().item() < 2e-2


def test_scheduler_metadata_mismatch_is_clamped_on_device():
    """A schedule from get_scheduler_metadata built for one plan must not make another plan write out of bounds: the per-
    batch split count is clamped to the partials the call allocated, and a header that does not match the plan (other
    num_splits / pack_gqa / sm_margin) is refused by the kernel -- covered here by the matching case staying correct when
    the schedule asks for more splits than a later call with the SAME plan but shorter sequences needs."""
    F = _iface()
    g = torch.Generator().manual_seed(64)
    b, h, hk, d, cap = 4, 8, 2, 128, 4096
    q = torch.randn(b, 1, h, d, generator=g).to(torch.bfloat16)
    kc, vc = torch.randn(b, cap, hk, d, generator=g).to(torch.bfloat16), torch.randn(b, cap, hk, d, generator=g).to(torch.bfloat16)
    long_lens = torch.tensor([4096, 4096, 4000, 3900], dtype=torch.int32)
    short_lens = torch.tensor([10, 4096, 129, 1], dtype=torch.int32)
    meta = F.get_scheduler_metadata(b, 1, cap, h, hk, d, long_lens.to(DEV), qkv_dtype=torch.bfloat16, num_splits=8)
    for lens in (long_lens, short_lens):
        out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kc.to(DEV), vc.to(DEV), cache_seqlens=lens.to(DEV), scheduler_metadata=meta,
                                                 num_splits=8, return_softmax_lse=True)
        ref, lse_ref = A.attention_oracle(q, kc, vc, k_lens=lens)
        low, _ = A.attention_oracle(q, kc, vc, k_lens=lens, upcast=False, reorder_ops=True)
        _check(out, lse, ref, low, lse_ref)


# ----------------------------------------------------------------------------- qv
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("mha_type", ["mha", "mqa", "gqa"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("sq,sk", [(1, 1), (1, 1024), (3, 799), (64, 128), (128, 128), (113, 203), (300, 411), (1023, 1024)])
def test_qv(sq, sk, causal, mha_type, dtype):
    """qv: scores = (q k^T + qv v^T) * softmax_scale with d = 64, d_v = 256 and the default scale 1/sqrt(d + d_v)
    (hopper/flash_api.cpp:1028-1050, mainloop_fwd_sm90_tma_gmma_ws.hpp:1144-1148; the reference tests it with d = 64,
    d_v in {256, 512}: hopper/test_flash_attn.py:116-127).  Here V is read twice from its shared-memory stage: K-major for
    the score term, MN-major for P V."""
    F = _iface()
    h = 6
    hk = {"mha": 6, "mqa": 1, "gqa": 2}[mha_type]
    g = torch.Generator().manual_seed(71)
    mk = lambda *s: torch.randn(*s, generator=g).to(dtype)  # noqa: E731
    q, k, v, qv = mk(2, sq, h, 64), mk(2, sk, hk, 64), mk(2, sk, hk, 256), mk(2, sq, h, 256)
    ref, lse_ref = A.attention_oracle(q, k, v, qv=qv, causal=causal)
    low, _ = A.attention_oracle(q, k, v, qv=qv, causal=causal, upcast=False, reorder_ops=True)
    for pack_gqa, num_splits in ((False, 1), (True, 1), (None, 3)):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), qv=qv.to(DEV), causal=causal, pack_gqa=pack_gqa, num_splits=num_splits,
                                     return_attn_probs=True)
        _check(out, lse, ref, low, lse_ref)


def test_qv_varlen_and_paged_decode():
    """qv through the varlen entry point and through the paged KV cache with MLA-like head counts (16 query heads on one
    KV head, packed into the M dimension)."""
    F = _iface()
    g = torch.Generator().manual_seed(72)
    dt = torch.bfloat16
    lens_q, lens_k = [5, 200, 131], [64, 333, 131]
    cq = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), dtype=torch.int32)
    ck = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), dtype=torch.int32)
    h, hk = 4, 2
    q, qv = torch.randn(int(cq[-1]), h, 64, generator=g).to(dt), torch.randn(int(cq[-1]), h, 256, generator=g).to(dt)
    k, v = torch.randn(int(ck[-1]), hk, 64, generator=g).to(dt), torch.randn(int(ck[-1]), hk, 256, generator=g).to(dt)
    out, lse = F.flash_attn_varlen_func(q.to(DEV), k.to(DEV), v.to(DEV), cq.to(DEV), ck.to(DEV), max(lens_q), max(lens_k), qv=qv.to(DEV),
                                        causal=True, return_attn_probs=True)
    for i in range(3):
        qs, ks = slice(int(cq[i]), int(cq[i + 1])), slice(int(ck[i]), int(ck[i + 1]))
        ref, lse_ref = A.attention_oracle(q[qs][None], k[ks][None], v[ks][None], qv=qv[qs][None], causal=True)
        low, _ = A.attention_oracle(q[qs][None], k[ks][None], v[ks][None], qv=qv[qs][None], causal=True, upcast=False, reorder_ops=True)
        _check(out[qs][None], lse[:, qs][None], ref, low, lse_ref)
    # decode: b 3, 16 query heads / 1 KV head, paged cache, ragged lengths, library-chosen and forced splits
    b, h, hk, ps, cap = 3, 16, 1, 64, 1024
    q, qv = torch.randn(b, 1, h, 64, generator=g).to(dt), torch.randn(b, 1, h, 256, generator=g).to(dt)
    kd, vd, table, kp, vp = A.paged_kvcache(cap, ps, b, hk, 64, 256, dt, generator=g)
    lens = torch.tensor([1024, 517, 1], dtype=torch.int32)
    ref, lse_ref = A.attention_oracle(q, kd, vd, qv=qv, k_lens=lens)
    low, _ = A.attention_oracle(q, kd, vd, qv=qv, k_lens=lens, upcast=False, reorder_ops=True)
    for num_splits in (0, 1, 4):
        out, lse, *_ = F.flash_attn_with_kvcache(q.to(DEV), kp.to(DEV), vp.to(DEV), qv=qv.to(DEV), cache_seqlens=lens.to(DEV),
                                                 page_table=table.to(DEV), num_splits=num_splits, return_softmax_lse=True)
        _check(out, lse, ref, low, lse_ref)


def test_qv_rules():
    """The reference's argument rules for qv (hopper/flash_api.cpp:1028-1050) raise the same way."""
    F = _iface()
    mk = lambda *s, dt=torch.bfloat16: torch.zeros(*s, dtype=dt, device=DEV)  # noqa: E731
    with pytest.raises(RuntimeError, match="head_size <= 64"):
        F.flash_attn_func(mk(1, 8, 2, 128), mk(1, 8, 2, 128), mk(1, 8, 2, 256), qv=mk(1, 8, 2, 256))
    with pytest.raises(RuntimeError, match="hdim_v >= 256"):
        F.flash_attn_func(mk(1, 8, 2, 64), mk(1, 8, 2, 64), mk(1, 8, 2, 128), qv=mk(1, 8, 2, 128))
    with pytest.raises(RuntimeError, match="wrong shape"):
        F.flash_attn_func(mk(1, 8, 2, 64), mk(1, 8, 2, 64), mk(1, 8, 2, 256), qv=mk(1, 9, 2, 256))


@pytest.mark.parametrize("dtype", [torch.bfloat16, E4M3])
@pytest.mark.parametrize("kw", [dict(causal=True), dict(window_size=(300, 0)), dict(window_size=(200, 100)), dict(causal=True, softcap=20.0)])
@pytest.mark.parametrize("b,sq,sk,h,hk,d", [(3, 1500, 1500, 8, 2, 128), (2, 1100, 2300, 6, 6, 64), (5, 700, 700, 4, 1, 128), (2, 900, 900, 4, 2, 256)])
def test_dynamic_schedule(b, sq, sk, h, hk, d, kw, dtype, monkeypatch):
    """Dynamic tile scheduling (Fa100FwdParams.tile_counter, the reference's tile_count_semaphore): the CTAs take the items of the
    length-ordered list as they finish.  Which CTA computes an item does not change its arithmetic: the result is bitwise the one
    of the static schedule, it matches the oracle, and a second call (fresh counter) repeats it."""
    import fa100_ops

    F = _iface()
    fp8 = dtype == E4M3
    g = torch.Generator().manual_seed(21)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16).to(dtype)  # noqa: E731
    q, k, v = mk(b, sq, h, d), mk(b, sk, hk, d), mk(b, sk, hk, d)
    qr, kr, vr = (x.to(torch.bfloat16) if fp8 else x for x in (q, k, v))
    ref, lse_ref = A.attention_oracle(qr, kr, vr, **kw)
    low, _ = A.attention_oracle(qr, kr, vr, upcast=False, reorder_ops=True, intermediate_dtype=dtype if fp8 else None, **kw)
    monkeypatch.setattr(fa100_ops, "_DYNAMIC", False)
    static, _ = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), return_attn_probs=True, **kw)
    monkeypatch.setattr(fa100_ops, "_DYNAMIC", True)
    for num_splits in (1, 3):
        out, lse = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), num_splits=num_splits, return_attn_probs=True, **kw)
        _check(out, lse, ref, low, lse_ref, rtol=4.0 if fp8 else (3.0 if kw.get("softcap") else 2.0))
        if num_splits == 1:
            assert torch.equal(out, static)
            again, _ = F.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), return_attn_probs=True, **kw)
            assert torch.equal(out, again)


def test_dynamic_schedule_cuda_graph():
    """The tile counter is allocated and zeroed inside the op (torch.zeros): under CUDA-graph capture that is a memset node in front
    of the kernel, so every replay starts from zero and reproduces the eager result bitwise."""
    F = _iface()
    g = torch.Generator().manual_seed(22)
    q = torch.randn(2, 1300, 8, 128, generator=g).to(torch.bfloat16).to(DEV)
    k = torch.randn(2, 1300, 2, 128, generator=g).to(torch.bfloat16).to(DEV)
    v = torch.randn(2, 1300, 2, 128, generator=g).to(torch.bfloat16).to(DEV)
    eager = F.flash_attn_func(q, k, v, causal=True)
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        F.flash_attn_func(q, k, v, causal=True)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = F.flash_attn_func(q, k, v, causal=True)
    torch.cuda.current_stream().wait_stream(st)
    for _ in range(3):
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager)


@pytest.mark.parametrize("causal", [False, True])
def test_tma_store_stays_inside_sequences(causal):
    """Packed and seqused_q batches take the TMA-store epilogue only for tiles that lie wholly inside their sequence (o_tma_check):
    rows past a sequence's used length keep what the caller's buffer held, and no sequence's first rows are overwritten by the box of
    the sequence in front of it (lengths that are not multiples of the 128-row tile)."""
    import fa100_ops

    g = torch.Generator().manual_seed(31)
    h, hk, d = 4, 2, 128
    # dense layout + seqused_q: rows [used, seqlen_q) must keep the sentinel
    b, sq, sk = 3, 700, 900
    q = torch.randn(b, sq, h, d, generator=g).to(torch.bfloat16)
    k = torch.randn(b, sk, hk, d, generator=g).to(torch.bfloat16)
    v = torch.randn(b, sk, hk, d, generator=g).to(torch.bfloat16)
    used = torch.tensor([300, 513, 700], dtype=torch.int32)
    out = torch.full((b, sq, h, d), 7.0, dtype=torch.bfloat16, device=DEV)
    for num_splits in (1, 0, 3):  # one pass; the library's choice (split-KV here: few tiles) and a forced split, both through fa100_combine
      out.fill_(7.0)
      res = fa100_ops.mha_fwd(q.to(DEV), k.to(DEV), v.to(DEV), out=out, seqused_q=used.to(DEV), is_causal=causal, num_splits=num_splits)
      torch.cuda.synchronize()
      o = res[0].cpu()
      for i in range(b):
          u = int(used[i])
          assert torch.all(o[i, u:] == 7.0), "rows past seqused_q were written"
          ref, _ = A.attention_oracle(q[i:i + 1, :u], k[i:i + 1], v[i:i + 1], causal=causal)
          low, _ = A.attention_oracle(q[i:i + 1, :u], k[i:i + 1], v[i:i + 1], causal=causal, upcast=False, reorder_ops=True)
          _check(o[i:i + 1, :u], None, ref, low, None)
    # packed layout: every sequence against the oracle (a box that ran over would have replaced the next sequence's first rows)
    F = _iface()
    lens = [130, 257, 1, 640, 383, 128, 900]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    tot = int(cu[-1])
    q = torch.randn(tot, h, d, generator=g).to(torch.bfloat16)
    k = torch.randn(tot, hk, d, generator=g).to(torch.bfloat16)
    v = torch.randn(tot, hk, d, generator=g).to(torch.bfloat16)
    for _ in range(3):  # the order in which the CTAs finish varies from launch to launch
        o, lse = F.flash_attn_varlen_func(q.to(DEV), k.to(DEV), v.to(DEV), cu.to(DEV), cu.to(DEV), max(lens), max(lens), causal=causal,
                                          return_attn_probs=True)
        for i, n in enumerate(lens):
            sl = slice(int(cu[i]), int(cu[i + 1]))
            ref, lse_ref = A.attention_oracle(q[sl][None], k[sl][None], v[sl][None], causal=causal)
            low, _ = A.attention_oracle(q[sl][None], k[sl][None], v[sl][None], causal=causal, upcast=False, reorder_ops=True)
            _check(o[sl][None], lse[:, sl][None], ref, low, lse_ref)
