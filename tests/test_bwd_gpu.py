"""Backward pass (torch.ops.flash_attn_3.bwd -> fa100_bwd) against gradients of the CPU oracle.

Rule of hopper/test_flash_attn.py:225-286: for g in (dq, dk, dv)
    max|g - g_ref| <= 2 * max|g_pt - g_ref| + 2 * max|(g_ref + 0.3 - 0.3) - g_ref|
with g_ref from the fp32 oracle and g_pt from its low-precision arm (upcast=False, reorder_ops=True), all by torch autograd
on the CPU.  Scope: bf16 / fp16, d == dv <= 128, MHA / GQA / MQA, fixed-length and variable-length batches, causal, local windows,
softcap."""
import pytest
import torch

import attention_oracle as A

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _grads(q, k, v, g, **kw):
    q, k, v = (x.detach().clone().requires_grad_() for x in (q, k, v))
    out, _ = A.attention_oracle(q, k, v, **kw)
    return torch.autograd.grad(out, (q, k, v), g)


def _check_grads(got, ref, pt, names=("dq", "dk", "dv")):
    for a, r, p_, nm in zip(got, ref, pt, names):
        assert torch.isfinite(a.float()).all(), nm
        atol = 2 * (r + 0.3 - 0.3 - r).abs().max().item()
        err = (a.float().cpu() - r.float()).abs().max().item()
        # + an absolute floor for gradients that are exactly zero in both oracle arms (a single key: P = 1, dS = 0): there the
        # rule's right-hand side is 0, while D (summed by the preprocess kernel) and dP (summed by the MMA) differ by fp32
        # summation order, ~1e-6 relative, accumulated over the rows; gradients elsewhere are O(1..50)
        tol = 2 * (p_.float() - r.float()).abs().max().item() + atol + 1e-4
        assert err <= tol, f"{nm}: max err {err:.3e} > tol {tol:.3e}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("mha_type", ["mha", "mqa", "gqa"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128])
@pytest.mark.parametrize("sq,sk", [(1, 1), (64, 128), (128, 128), (113, 203), (128, 217), (108, 256), (256, 512), (239, 1), (384, 256),
                                   (640, 128), (512, 256), (1024, 1024), (1023, 1024), (799, 3), (1, 1024)])
def test_backward(sq, sk, d, causal, mha_type, dtype):
    import flash_attn_interface as F

    h = 6
    hk = {"mha": 6, "mqa": 1, "gqa": 2}[mha_type]
    gen = torch.Generator().manual_seed(101)
    q, k, v = (torch.randn(2, s_, h_, d, generator=gen).to(dtype) for s_, h_ in ((sq, h), (sk, hk), (sk, hk)))
    g = torch.randn(2, sq, h, d, generator=gen).to(dtype)
    ref = _grads(q.float(), k.float(), v.float(), g.float(), causal=causal)
    pt = _grads(q, k, v, g, causal=causal, upcast=False, reorder_ops=True)
    qd, kd, vd = (x.to(DEV).requires_grad_() for x in (q, k, v))
    out = F.flash_attn_func(qd, kd, vd, causal=causal)
    dq, dk, dv = torch.autograd.grad(out, (qd, kd, vd), g.to(DEV))
    torch.cuda.synchronize()
    _check_grads((dq, dk, dv), ref, pt)


@pytest.mark.parametrize("d", [40, 72, 96])
def test_backward_odd_head_dims(d):
    import flash_attn_interface as F

    gen = torch.Generator().manual_seed(102)
    q, k, v = (torch.randn(2, s_, h_, d, generator=gen).to(torch.bfloat16) for s_, h_ in ((300, 4), (411, 2), (411, 2)))
    g = torch.randn(2, 300, 4, d, generator=gen).to(torch.bfloat16)
    ref = _grads(q.float(), k.float(), v.float(), g.float(), causal=True)
    pt = _grads(q, k, v, g, causal=True, upcast=False, reorder_ops=True)
    qd, kd, vd = (x.to(DEV).requires_grad_() for x in (q, k, v))
    dq, dk, dv = torch.autograd.grad(F.flash_attn_func(qd, kd, vd, causal=True), (qd, kd, vd), g.to(DEV))
    _check_grads((dq, dk, dv), ref, pt)


def test_backward_op_outputs_and_determinism():
    """The op returns the reference's 8-tuple; softmax_d = rowsum(dout * out) (hopper/test_flash_attn.py:234,248-249); two
    runs are bitwise identical (no atomics); strided (non-contiguous) inputs work."""
    import flash_attn_interface as F  # noqa: F401  (registers the ops)

    gen = torch.Generator(device=DEV).manual_seed(103)
    b, s, h, hk, d = 2, 500, 8, 2, 128
    qkv = torch.randn(b, s, h + 2 * hk, d, generator=gen, device=DEV, dtype=torch.bfloat16)
    q, k, v = qkv.split([h, hk, hk], dim=2)  # views with a head stride of the packed tensor
    out, lse, *_ = torch.ops.flash_attn_3.fwd(q, k, v, is_causal=True)
    g = torch.randn(out.shape, generator=gen, device=DEV, dtype=torch.bfloat16)
    res = torch.ops.flash_attn_3.bwd(g, q, k, v, out, lse, None, None, None, None, None, None, None, None, None, None, True)
    assert len(res) == 8
    dq, dk, dv, softmax_d = res[:4]
    do_o = (g.float() * out.float()).sum(-1).transpose(1, 2)
    assert torch.allclose(softmax_d[:, :, :s], do_o, atol=1e-3, rtol=1e-3) and softmax_d.shape[-1] == 512
    res2 = torch.ops.flash_attn_3.bwd(g, q, k, v, out, lse, None, None, None, None, None, None, None, None, None, None, True)
    assert all(torch.equal(a, b_) for a, b_ in zip(res[:3], res2[:3]))
    ref = _grads(q.cpu().float(), k.cpu().float(), v.cpu().float(), g.cpu().float(), causal=True)
    pt = _grads(q.cpu(), k.cpu(), v.cpu(), g.cpu(), causal=True, upcast=False, reorder_ops=True)
    _check_grads((dq, dk, dv), ref, pt)


@pytest.mark.parametrize("kw", [dict(window_size=(100, 50)), dict(window_size=(300, 0)), dict(window_size=(0, 0)), dict(window_size=(-1, 30)),
                                dict(window_size=(20, -1)), dict(softcap=15.0), dict(softcap=15.0, causal=True), dict(softcap=30.0, window_size=(64, 16))])
@pytest.mark.parametrize("sq,sk", [(500, 700), (700, 500), (64, 1500), (1, 333), (256, 256)])
def test_backward_local_and_softcap(sq, sk, kw):
    """Sliding windows and softcap in the backward (hopper/test_flash_attn.py:63,62 x :225-286; rtol 3 with softcap, :195)."""
    import flash_attn_interface as F

    gen = torch.Generator().manual_seed(104)
    q, k, v = (torch.randn(2, s_, h_, 128, generator=gen).to(torch.bfloat16) for s_, h_ in ((sq, 4), (sk, 2), (sk, 2)))
    g = torch.randn(2, sq, 4, 128, generator=gen).to(torch.bfloat16)
    ref = _grads(q.float(), k.float(), v.float(), g.float(), **kw)
    pt = _grads(q, k, v, g, upcast=False, reorder_ops=True, **kw)
    qd, kd, vd = (x.to(DEV).requires_grad_() for x in (q, k, v))
    dq, dk, dv = torch.autograd.grad(F.flash_attn_func(qd, kd, vd, **kw), (qd, kd, vd), g.to(DEV))
    torch.cuda.synchronize()
    rt = 3.0 if kw.get("softcap") else 2.0
    for a_, r, p_, nm in zip((dq, dk, dv), ref, pt, ("dq", "dk", "dv")):
        atol = 2 * (r + 0.3 - 0.3 - r).abs().max().item() + (3e-4 if kw.get("softcap") else 0)
        err = (a_.float().cpu() - r.float()).abs().max().item()
        assert torch.isfinite(a_.float()).all() and err <= rt * (p_.float() - r.float()).abs().max().item() + atol + 1e-4, (nm, err)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("use_seqused", [False, True])
@pytest.mark.parametrize("lens_q,lens_k", [([100, 257, 31], [100, 300, 77]), ([128, 5, 300, 1], [128, 64, 511, 1000]),
                                           ([1, 1, 1, 1, 1], [7, 130, 250, 1, 600])])
def test_backward_varlen(lens_q, lens_k, use_seqused, causal):
    """Packed variable-length batches (cu_seqlens), optionally with seqused shorter than the slots (hopper/test_flash_attn.py:290-598)."""
    import flash_attn_interface as F

    gen = torch.Generator().manual_seed(105)
    h, hk, d = 4, 2, 128
    b = len(lens_q)
    slot_q = [n + (17 if use_seqused else 0) for n in lens_q]   # with seqused every slot holds unused rows after the sequence
    slot_k = [n + (40 if use_seqused else 0) for n in lens_k]
    cq = torch.tensor([0] + torch.tensor(slot_q).cumsum(0).tolist(), dtype=torch.int32)
    ck = torch.tensor([0] + torch.tensor(slot_k).cumsum(0).tolist(), dtype=torch.int32)
    q = torch.randn(int(cq[-1]), h, d, generator=gen).to(torch.bfloat16)
    k = torch.randn(int(ck[-1]), hk, d, generator=gen).to(torch.bfloat16)
    v = torch.randn(int(ck[-1]), hk, d, generator=gen).to(torch.bfloat16)
    g = torch.randn(int(cq[-1]), h, d, generator=gen).to(torch.bfloat16)
    qd, kd, vd = (x.to(DEV).requires_grad_() for x in (q, k, v))
    extra = dict(seqused_q=torch.tensor(lens_q, dtype=torch.int32, device=DEV), seqused_k=torch.tensor(lens_k, dtype=torch.int32, device=DEV)) if use_seqused else {}
    out = F.flash_attn_varlen_func(qd, kd, vd, cq.to(DEV), ck.to(DEV), max(slot_q), max(slot_k), causal=causal, **extra)
    dq, dk, dv = torch.autograd.grad(out, (qd, kd, vd), g.to(DEV))
    torch.cuda.synchronize()
    for i in range(b):
        qs, ks = slice(int(cq[i]), int(cq[i]) + lens_q[i]), slice(int(ck[i]), int(ck[i]) + lens_k[i])
        ref = _grads(q[qs][None].float(), k[ks][None].float(), v[ks][None].float(), g[qs][None].float(), causal=causal)
        pt = _grads(q[qs][None], k[ks][None], v[ks][None], g[qs][None], causal=causal, upcast=False, reorder_ops=True)
        _check_grads((dq[qs][None], dk[ks][None], dv[ks][None]), ref, pt)
        if use_seqused:  # the unused tail of every slot gets zero gradients
            assert dq[int(cq[i]) + lens_q[i]:int(cq[i + 1])].abs().max().item() == 0
            assert dk[int(ck[i]) + lens_k[i]:int(ck[i + 1])].abs().max().item() == 0 and dv[int(ck[i]) + lens_k[i]:int(ck[i + 1])].abs().max().item() == 0


def test_backward_unsupported_raises():
    import flash_attn_interface as F

    mk = lambda *s: torch.zeros(*s, dtype=torch.bfloat16, device=DEV, requires_grad=True)  # noqa: E731
    q2, k2, v2 = mk(1, 64, 2, 256), mk(1, 64, 2, 256), mk(1, 64, 2, 256)
    with pytest.raises(NotImplementedError):
        F.flash_attn_func(q2, k2, v2).sum().backward()  # head dim > 128: not in this version
    q, k, v = mk(1, 64, 2, 128), mk(1, 64, 2, 128), mk(1, 64, 2, 128)
    with pytest.raises(AssertionError):
        F.flash_attn_func(q, k, v, attention_chunk=16, causal=True).sum().backward()  # as in the reference (flash_attn_interface.py:238)
