"""Packed variable-length prefill against the same tokens as a fixed-length batch.  python tools/varlen_bench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
from flash_attn_interface import flash_attn_func, flash_attn_varlen_func

def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

h, hk, d = 16, 16, 128
for name, lens in (("16 x 2048 equal", [2048] * 16), ("8 x 4096 equal", [4096] * 8), ("ragged 512..8192", [8192, 6144, 4096, 4096, 3072, 2048, 2048, 1024, 1024, 512, 512]),
                   ("64 x 512 equal", [512] * 64)):
    tot = sum(lens)
    q = torch.randn(tot, h, d, device="cuda", dtype=torch.bfloat16); k = torch.randn(tot, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
    for causal in (True, False):
        fl = sum(2 * h * s * (s / 2 if causal else s) * 2 * d for s in lens)
        ms = t(lambda: flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), causal=causal))
        line = f"{name:18s} causal={causal!s:5}: varlen {ms*1e3:7.1f} us {fl/ms/1e9:6.0f} TF"
        if len(set(lens)) == 1:
            b, s = len(lens), lens[0]
            ms2 = t(lambda: flash_attn_func(q.view(b, s, h, d), k.view(b, s, hk, d), v.view(b, s, hk, d), causal=causal))
            line += f" | fixed {ms2*1e3:7.1f} us {fl/ms2/1e9:6.0f} TF"
        print(line, flush=True)
