"""CPU: the oracle restatement against golden vectors produced by the reference's own
attention_ref / attention_combine_ref (tests/golden/make_golden.py)."""
import os

import pytest
import torch

import attention_oracle as A

GOLD = torch.load(os.path.join(os.path.dirname(__file__), "golden", "attention_ref.pt"), weights_only=False)
ATTN_CASES = [k for k in GOLD if not k.startswith("combine")]


@pytest.mark.parametrize("name", ATTN_CASES)
def test_attention_oracle_matches_reference(name):
    rec = GOLD[name]
    kw = dict(rec["kwargs"])
    o, lse = A.attention_oracle(rec["q"], rec["k"], rec["v"], q_lens=rec.get("q_lens"), k_lens=rec.get("k_lens"), **kw)
    # identical arithmetic on the same CPU: bit-exact for fp32 upcast, and for the low-precision arm
    assert o.dtype == rec["out"].dtype
    assert torch.equal(o, rec["out"]), (o.float() - rec["out"].float()).abs().max()
    # LSE is consistent with the reference probabilities: dead rows (rowsum 0) <-> lse = -inf
    dead = rec["attn_rowsum"] < 0.5
    ql = rec.get("q_lens")
    if ql is not None:
        valid = (torch.arange(o.shape[1]).view(1, -1) < ql.view(-1, 1)).view(o.shape[0], 1, -1)
        dead = dead | ~valid
    assert torch.equal(torch.isneginf(lse), dead)


def test_lse_definition():
    torch.manual_seed(0)
    q, k, v = torch.randn(2, 19, 4, 32), torch.randn(2, 23, 2, 32), torch.randn(2, 23, 2, 32)
    o, lse = A.attention_oracle(q, k, v)
    s = torch.einsum("bthd,bshd->bhts", q, k.repeat_interleave(2, 2)) / 32**0.5
    assert torch.allclose(lse, torch.logsumexp(s, -1), atol=1e-5)
    p = torch.exp(s - lse.unsqueeze(-1))
    assert torch.allclose(o, torch.einsum("bhts,bshd->bthd", p, v.repeat_interleave(2, 2)), atol=1e-5)


@pytest.mark.parametrize("name", ["combine_a", "combine_b"])
def test_combine_oracle_matches_reference(name):
    rec = GOLD[name]
    o, lse = A.combine_oracle(rec["out_partial"], rec["lse_partial"])
    assert torch.equal(o, rec["out"]) and torch.equal(lse, rec["lse"])


def test_sliced_oracle_equals_full():
    torch.manual_seed(1)
    q, k, v = torch.randn(2, 70, 4, 32), torch.randn(2, 90, 2, 32), torch.randn(2, 90, 2, 32)
    for causal in (False, True):
        o, l = A.attention_oracle(q, k, v, causal=causal)
        o2, l2 = A.attention_oracle_sliced(q, k, v, rows_per_call=16, causal=causal)
        assert torch.allclose(o, o2, atol=1e-5) and torch.allclose(l, l2, atol=1e-5)


def test_split_then_combine_is_identity():
    """size-independent property: attention over key halves merged by the combine math == whole."""
    torch.manual_seed(2)
    q, k, v = torch.randn(1, 33, 2, 32), torch.randn(1, 64, 2, 32), torch.randn(1, 64, 2, 32)
    o, lse = A.attention_oracle(q, k, v)
    parts = [A.attention_oracle(q, k[:, a:b], v[:, a:b]) for a, b in ((0, 20), (20, 64))]
    op = torch.stack([p[0] for p in parts])
    lp = torch.stack([p[1].transpose(1, 2) for p in parts])
    oc, lc = A.combine_oracle(op, lp)
    assert torch.allclose(oc, o, atol=1e-5) and torch.allclose(lc.transpose(1, 2), lse, atol=1e-5)


def test_rotary_oracle_norm_and_identity():
    torch.manual_seed(3)
    x = torch.randn(2, 5, 3, 32)
    ang = torch.rand(16, 8) * 6.28
    for inter in (False, True):
        y = A.rotary_oracle(x, ang.cos(), ang.sin(), seqlen_offsets=torch.tensor([0, 7]), interleaved=inter)
        assert torch.allclose(y.norm(dim=-1), x.norm(dim=-1), atol=1e-5)
        assert torch.equal(y[..., 16:], x[..., 16:])
    z = A.rotary_oracle(x, torch.ones(16, 8), torch.zeros(16, 8))
    assert torch.allclose(z, x)
