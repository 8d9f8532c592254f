"""One backward call on a short-sequence shape (for ncu).  python tools/bwd_one_short.py [causal 0/1]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
import flash_attn_interface as F  # noqa: F401
causal = bool(int(sys.argv[1])) if len(sys.argv) > 1 else False
b, s, h, hk = 32, 1024, 16, 16
q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16)
k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16)
v = torch.randn_like(k)
out, lse, *_ = torch.ops.flash_attn_3.fwd(q, k, v, is_causal=causal)
g = torch.randn_like(out)
for _ in range(3):
    torch.ops.flash_attn_3.bwd(g, q, k, v, out, lse, None, None, None, None, None, None, None, None, None, None, causal)
torch.cuda.synchronize()
print("ok")
