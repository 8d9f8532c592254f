"""Golden vectors for the rotary oracle from the REFERENCE's own pure-PyTorch implementation,
/root/reference/flash_attn/layers/rotary.py:14-36 (rotate_half, apply_rotary_emb_torch).  Run in the build container only:

    python tests/golden/make_golden_rotary.py

The module imports a Triton kernel at import time (flash_attn.ops.triton.rotary), so the two functions are taken from its
source text (ast) and executed unmodified.  Positions: the reference function takes per-batch (b, s, rotary_dim/2) cos / sin,
so ``seqlen_offsets`` become a gather of the table rows offset[b] + t -- what hopper/rotary.h:19-77 does with seqlens_rotary.
Stores inputs + fp32 outputs in tests/golden/rotary_ref.pt; tests/test_oracle.py checks oracle/attention_oracle.rotary_oracle
against them."""
import ast
import os

import torch
from einops import rearrange, repeat  # noqa: F401  (used by the extracted functions)

SRC = "/root/reference/flash_attn/layers/rotary.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_functions():
    tree = ast.parse(open(SRC).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("rotate_half", "apply_rotary_emb_torch")]
    assert len(keep) == 2
    ns = dict(torch=torch, rearrange=rearrange, repeat=repeat)
    exec(compile(ast.Module(body=keep, type_ignores=[]), SRC, "exec"), ns)
    return ns["apply_rotary_emb_torch"]


def main():
    ref = load_reference_functions()
    out = {}
    cases = [("full_halves", 3, 7, 4, 64, 64, False), ("full_interleaved", 3, 7, 4, 64, 64, True), ("partial_halves", 2, 5, 2, 128, 64, False),
             ("partial_interleaved", 2, 5, 2, 128, 32, True), ("one_token", 4, 1, 6, 128, 128, False), ("rd16", 2, 9, 2, 64, 16, True)]
    for i, (name, b, s, h, d, rd, inter) in enumerate(cases):
        g = torch.Generator().manual_seed(2000 + i)
        x = torch.randn(b, s, h, d, generator=g).to(torch.bfloat16)
        ang = torch.rand(64, rd // 2, generator=g) * 6.283185307179586
        cos, sin = torch.cos(ang).to(torch.bfloat16), torch.sin(ang).to(torch.bfloat16)
        off = torch.randint(0, 64 - s, (b,), generator=g)
        idx = off.view(b, 1) + torch.arange(s).view(1, s)
        y = ref(x.float(), cos.float()[idx], sin.float()[idx], interleaved=inter)
        out[name] = dict(x=x, cos=cos, sin=sin, offsets=off, interleaved=inter, out_fp32=y)
    torch.save(out, os.path.join(HERE, "rotary_ref.pt"))
    print("wrote", len(out), "cases")


if __name__ == "__main__":
    main()
