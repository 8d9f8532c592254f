"""Backward timing on the C2 shape and an MHA shape (TFLOP/s counted as 2.5 x the forward's algorithmic FLOPs, the convention of
hopper/benchmark_attn.py:62-73).  python tools/bwd_bench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
import flash_attn_interface as F  # noqa: F401
for (b, s, h, hk, causal) in ((8, 8192, 32, 8, True), (4, 8192, 16, 16, False), (2, 16384, 16, 16, True), (32, 1024, 16, 16, True), (16, 2048, 16, 16, True),
                            (32, 1024, 16, 16, False), (8, 4096, 16, 16, True)):
    q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16)
    v = torch.randn_like(k)
    out, lse, *_ = torch.ops.flash_attn_3.fwd(q, k, v, is_causal=causal)
    g = torch.randn_like(out)
    f = lambda: torch.ops.flash_attn_3.bwd(g, q, k, v, out, lse, None, None, None, None, None, None, None, None, None, None, causal)
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); [f() for _ in range(5)]; e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    fl = 2.5 * b * h * 2 * s * (s / 2 if causal else s) * 256
    print(f"bwd b{b} s{s} h{h}/{hk} causal={causal}: {ms:.2f} ms, {fl/ms/1e9:.0f} TFLOP/s", flush=True)
