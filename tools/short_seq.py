"""Short-sequence efficiency: eager vs CUDA-graph timing, batch scaling.  python tools/short_seq.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
from flash_attn_interface import flash_attn_func

def run(b, s, h, hk, d, causal, graph):
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k)
    f = lambda: flash_attn_func(q, k, v, causal=causal)
    for _ in range(3): f()
    torch.cuda.synchronize()
    n = 20
    if graph:
        g = torch.cuda.CUDAGraph()
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            f(); torch.cuda.synchronize()
            with torch.cuda.graph(g):
                for _ in range(n): f()
        torch.cuda.synchronize()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
    else:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): f()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
    fl = b * h * 2 * s * (s / 2 if causal else s) * 2 * d
    return ms, fl / ms / 1e9

for (b, s, h, hk, d, causal) in ((32, 1024, 16, 16, 128, True), (128, 1024, 16, 16, 128, True), (32, 1024, 16, 16, 128, False), (16, 2048, 16, 16, 128, True),
                                 (32, 1024, 32, 32, 64, True), (8, 4096, 16, 16, 128, True), (4, 8192, 16, 16, 128, True), (16, 8192, 16, 16, 128, True)):
    a = run(b, s, h, hk, d, causal, False); g = run(b, s, h, hk, d, causal, True)
    print(f"b{b} s{s} h{h}/{hk} d{d} causal={causal}: eager {a[0]*1e3:7.1f} us {a[1]:6.0f} TF | graph {g[0]*1e3:7.1f} us {g[1]:6.0f} TF", flush=True)
