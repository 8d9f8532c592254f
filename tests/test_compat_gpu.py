"""The FA2-style entry points (flash-attention-annotated_b200/compat/flash_attn/flash_attn_interface.py)
against the CPU oracle, and -- where the FA2 wheel of the image is importable -- against that library's own kernels on the
same inputs (a cross-check only: library code, never on the product path).  Signatures are pinned on CPU by
tests/test_abi_cpu.py::test_fa2_compat_signatures_match_reference."""
import importlib.util
import math
import os

import pytest
import torch

import attention_oracle as A

pytestmark = pytest.mark.gpu
DEV = "cuda"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _shim():
    path = os.path.join(ROOT, "flash-attention-annotated_b200", "compat", "flash_attn", "flash_attn_interface.py")
    spec = importlib.util.spec_from_file_location("fa100_compat_flash_attn_interface", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _fa2_wheel():
    try:
        import flash_attn  # the FA2 wheel in site-packages (the compat directory is not on sys.path in the tests)

        return flash_attn if "compat" not in (flash_attn.__file__ or "") else None
    except Exception:  # noqa: BLE001
        return None


def _close(out, ref, low, rtol=2.0):
    torch.cuda.synchronize()
    err = (out.float().cpu() - ref.float()).abs().max().item()
    assert err <= A.fwd_tolerance(ref, low, rtol), err


@pytest.mark.parametrize("d", [59, 64, 96, 128])
@pytest.mark.parametrize("kw", [dict(), dict(causal=True), dict(window_size=(64, 0)), dict(softcap=30.0, causal=True)])
def test_compat_flash_attn_func(d, kw):
    C = _shim()
    g = torch.Generator().manual_seed(81)
    q, k, v = (torch.randn(2, s_, h_, d, generator=g).to(torch.bfloat16) for s_, h_ in ((200, 6), (333, 2), (333, 2)))
    ref, lse_ref = A.attention_oracle(q, k, v, **kw)
    low, _ = A.attention_oracle(q, k, v, upcast=False, reorder_ops=True, **kw)
    out = C.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), 0.0, **kw)
    assert out.shape == q.shape
    _close(out, ref, low, 3.0 if kw.get("softcap") else 2.0)
    out2, lse, s_dmask = C.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), return_attn_probs=True, **kw)
    assert s_dmask is None and torch.equal(out2, out)
    assert torch.allclose(lse.cpu(), lse_ref, atol=1e-3, rtol=1e-3)
    fa2 = _fa2_wheel()
    if fa2 is not None:  # the FA2 library's own kernel on the same inputs
        theirs = fa2.flash_attn_func(q.to(DEV), k.to(DEV), v.to(DEV), 0.0, **kw)
        assert (out.float() - theirs.float()).abs().max().item() <= 2e-2


def test_compat_packed_variants_and_unsupported():
    C = _shim()
    g = torch.Generator().manual_seed(82)
    qkv = torch.randn(2, 150, 3, 4, 64, generator=g).to(torch.bfloat16).to(DEV)
    base = C.flash_attn_func(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=True)
    assert torch.equal(C.flash_attn_qkvpacked_func(qkv, causal=True), base)
    assert torch.equal(C.flash_attn_kvpacked_func(qkv[:, :, 0], qkv[:, :, 1:], causal=True), base)
    cu = torch.tensor([0, 150, 300], dtype=torch.int32, device=DEV)
    packed = qkv.reshape(300, 3, 4, 64)
    v1 = C.flash_attn_varlen_qkvpacked_func(packed, cu, 150, causal=True)
    v2 = C.flash_attn_varlen_kvpacked_func(packed[:, 0], packed[:, 1:], cu, cu, 150, 150, causal=True)
    assert torch.equal(v1, v2) and (v1.view(2, 150, 4, 64).float() - base.float()).abs().max().item() <= 1e-2
    with pytest.raises(NotImplementedError):
        C.flash_attn_func(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], dropout_p=0.1)
    with pytest.raises(NotImplementedError):
        C.flash_attn_func(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], alibi_slopes=torch.ones(4, device=DEV))


@pytest.mark.parametrize("causal", [False, True])
def test_compat_varlen_block_table(causal):
    """flash_attn_varlen_func with a paged K/V cache addressed by block_table (flash_attn_interface.py:1386-1392)."""
    C = _shim()
    g = torch.Generator().manual_seed(83)
    b, h, hk, d, ps, cap = 3, 8, 2, 128, 256, 1024
    lens_q, lens_k = [5, 64, 1], [700, 1024, 33]
    cq = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), dtype=torch.int32)
    ck = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), dtype=torch.int32)
    q = torch.randn(int(cq[-1]), h, d, generator=g).to(torch.bfloat16)
    kd, vd, table, kp, vp = A.paged_kvcache(cap, ps, b, hk, d, d, torch.bfloat16, generator=g)
    out = C.flash_attn_varlen_func(q.to(DEV), kp.to(DEV), vp.to(DEV), cq.to(DEV), ck.to(DEV), max(lens_q), max(lens_k), causal=causal,
                                   block_table=table.to(DEV))
    for i in range(b):
        qs = slice(int(cq[i]), int(cq[i + 1]))
        kl = torch.tensor([lens_k[i]], dtype=torch.int32)
        ref, _ = A.attention_oracle(q[qs][None], kd[i:i + 1], vd[i:i + 1], k_lens=kl, causal=causal)
        low, _ = A.attention_oracle(q[qs][None], kd[i:i + 1], vd[i:i + 1], k_lens=kl, causal=causal, upcast=False, reorder_ops=True)
        _close(out[qs][None], ref, low)


def test_compat_kvcache_int_seqlens_append_rotary_block_table():
    C = _shim()
    g = torch.Generator().manual_seed(84)
    b, h, hk, d, cap = 2, 8, 2, 128, 512
    dt = torch.bfloat16
    q = torch.randn(b, 1, h, d, generator=g).to(dt)
    kc, vc = torch.randn(b, cap, hk, d, generator=g).to(dt), torch.randn(b, cap, hk, d, generator=g).to(dt)
    kn, vn = torch.randn(b, 1, hk, d, generator=g).to(dt), torch.randn(b, 1, hk, d, generator=g).to(dt)
    ang = torch.rand(cap + 8, d // 2, generator=g) * 2 * math.pi
    cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)
    n0 = 300  # cache_seqlens given as a plain int
    lens = torch.full((b,), n0, dtype=torch.int32)
    kn_ref = A.rotary_oracle(kn, cos, sin, lens, interleaved=True)
    q_ref = A.rotary_oracle(q, cos, sin, lens, interleaved=True)
    kd, vd = kc.clone(), vc.clone()
    kd[:, n0:n0 + 1], vd[:, n0:n0 + 1] = kn_ref, vn
    ref, lse_ref = A.attention_oracle(q_ref, kd, vd, k_lens=lens + 1, causal=True)
    low, _ = A.attention_oracle(q_ref, kd, vd, k_lens=lens + 1, causal=True, upcast=False, reorder_ops=True)
    kc_dev, vc_dev = kc.to(DEV), vc.to(DEV)
    out, lse = C.flash_attn_with_kvcache(q.to(DEV), kc_dev, vc_dev, kn.to(DEV), vn.to(DEV), rotary_cos=cos.to(DEV), rotary_sin=sin.to(DEV),
                                         cache_seqlens=n0, causal=True, return_softmax_lse=True)
    _close(out, ref, low)
    assert torch.allclose(lse.cpu(), lse_ref, atol=1e-3, rtol=1e-3)
    assert torch.equal(vc_dev.cpu(), vd) and torch.allclose(kc_dev.cpu().float(), kd.float(), atol=1e-3, rtol=1e-3)
    fa2 = _fa2_wheel()
    if fa2 is not None:
        kc2, vc2 = kc.to(DEV), vc.to(DEV)
        theirs = fa2.flash_attn_with_kvcache(q.to(DEV), kc2, vc2, kn.to(DEV), vn.to(DEV), rotary_cos=cos.to(DEV), rotary_sin=sin.to(DEV),
                                             cache_seqlens=n0, causal=True)
        assert (out.float() - theirs.float()).abs().max().item() <= 2e-2
        assert torch.equal(vc2, vc_dev) and torch.allclose(kc2.float(), kc_dev.float(), atol=1e-2, rtol=1e-2)
    # paged: block_table
    kdp, vdp, table, kp, vp = A.paged_kvcache(cap, 128, b, hk, d, d, dt, generator=g)
    lens2 = torch.tensor([512, 77], dtype=torch.int32)
    out = C.flash_attn_with_kvcache(q.to(DEV), kp.to(DEV), vp.to(DEV), cache_seqlens=lens2.to(DEV), block_table=table.to(DEV))
    ref, _ = A.attention_oracle(q, kdp, vdp, k_lens=lens2)
    low, _ = A.attention_oracle(q, kdp, vdp, k_lens=lens2, upcast=False, reorder_ops=True)
    _close(out, ref, low)
