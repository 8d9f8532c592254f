"""One backward call on the C2 shape (for ncu captures of the backward kernels).  python tools/bwd_one.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
import flash_attn_interface as F  # noqa: F401
b, s, h, hk, causal = 8, 8192, 32, 8, True
q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16)
k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16)
v = torch.randn_like(k)
out, lse, *_ = torch.ops.flash_attn_3.fwd(q, k, v, is_causal=causal)
g = torch.randn_like(out)
for _ in range(2):
    torch.ops.flash_attn_3.bwd(g, q, k, v, out, lse, None, None, None, None, None, None, None, None, None, None, causal)
torch.cuda.synchronize()
print("ok")
