"""Backward pass (torch.ops.flash_attn_3.bwd -> fa100_bwd) against gradients of the CPU oracle.

Rule of hopper/test_flash_attn.py:225-286: for g in (dq, dk, dv)
    max|g - g_ref| <= 2 * max|g_pt - g_ref| + 2 * max|(g_ref + 0.3 - 0.3) - g_ref|
with g_ref from the fp32 oracle and g_pt from its low-precision arm (upcast=False, reorder_ops=True), all by torch autograd
on the CPU.  Scope of this first version: fixed-length batches, bf16 / fp16, d == dv <= 128, causal or full, MHA / GQA / MQA."""
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


def test_backward_unsupported_raises():
    import flash_attn_interface as F

    mk = lambda *s: torch.zeros(*s, dtype=torch.bfloat16, device=DEV, requires_grad=True)  # noqa: E731
    q, k, v = mk(1, 64, 2, 128), mk(1, 64, 2, 128), mk(1, 64, 2, 128)
    with pytest.raises(NotImplementedError):
        F.flash_attn_func(q, k, v, window_size=(8, 0)).sum().backward()
    with pytest.raises(NotImplementedError):
        F.flash_attn_func(q, k, v, softcap=10.0).sum().backward()
    q2, k2, v2 = mk(1, 64, 2, 256), mk(1, 64, 2, 256), mk(1, 64, 2, 256)
    with pytest.raises(NotImplementedError):
        F.flash_attn_func(q2, k2, v2).sum().backward()
