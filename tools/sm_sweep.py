"""Throughput per SM as a function of how many SMs the persistent grid uses (sm_margin): if per-SM throughput rises when
fewer SMs run, the full-size kernel is limited by something shared (board power, L2), not by its own instruction stream.
  python tools/sm_sweep.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
from flash_attn_interface import flash_attn_func
dev = "cuda"
nsm = torch.cuda.get_device_properties(0).multi_processor_count
for name, (b, s, h, hk, causal, dt) in (("bf16 C2", (8, 8192, 32, 8, True, torch.bfloat16)), ("bf16 nc", (4, 8192, 32, 8, False, torch.bfloat16)),
                                        ("fp8 C3", (4, 16384, 16, 16, True, torch.float8_e4m3fn)), ("fp8 C3nc", (4, 16384, 16, 16, False, torch.float8_e4m3fn))):
    q = torch.randn(b, s, h, 128, device=dev, dtype=torch.bfloat16).to(dt)
    k = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16).to(dt)
    v = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16).to(dt)
    for margin in (0, nsm // 4, nsm // 2, 3 * nsm // 4):
        for _ in range(2): flash_attn_func(q, k, v, causal=causal, sm_margin=margin)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10 if margin < nsm // 2 else 4
        e0.record()
        for _ in range(n): flash_attn_func(q, k, v, causal=causal, sm_margin=margin)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        fl = b * h * 2 * s * (s / 2 if causal else s) * 256
        tf = fl / ms / 1e9
        print(f"{name:9s} SMs {nsm - margin:3d}: {ms:8.3f} ms  {tf:7.1f} TFLOP/s  {tf / (nsm - margin):6.2f} per SM", flush=True)
