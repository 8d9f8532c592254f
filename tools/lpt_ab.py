"""Causal MHA / GQA shapes for the LPT super-group size A/B: FA100_LIB=... python tools/lpt_ab.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
from flash_attn_interface import flash_attn_func
res = []
for (b, s, h, hk) in ((4, 8192, 16, 16), (16, 8192, 16, 16), (2, 16384, 16, 16), (8, 4096, 16, 16), (8, 8192, 32, 8), (1, 32768, 32, 8)):
    q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16); k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k)
    for _ in range(3): flash_attn_func(q, k, v, causal=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): flash_attn_func(q, k, v, causal=True)
    e1.record(); torch.cuda.synchronize()
    res.append(b * h * 2 * s * (s / 2) * 256 / (e0.elapsed_time(e1) / 10) / 1e9)
print(os.environ.get("FA100_LIB", "libfa100.so"), " ".join(f"{x:7.0f}" for x in res))
