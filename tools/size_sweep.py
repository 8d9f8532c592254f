"""Cycles per key block of CTA 0 (cycle-bucket build) as the launch gets longer: same instruction stream, growing batch.
  FA100_LIB=libfa100_prof.so python tools/size_sweep.py [--fp8]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
os.environ.setdefault("FA100_LIB", "libfa100_prof.so")
import torch
import fa100_native as N
from flash_attn_interface import flash_attn_func
dt = torch.float8_e4m3fn if "--fp8" in sys.argv else torch.bfloat16
s, h, hk = 8192, 16, 4
causal = "--causal" in sys.argv
for b in (int(os.environ.get("SWEEP_B", "0")),) if os.environ.get("SWEEP_B") else (1, 2, 4, 8, 16, 32):
    q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16).to(dt)
    k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16).to(dt)
    v = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16).to(dt)
    flash_attn_func(q, k, v, causal=causal); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); flash_attn_func(q, k, v, causal=causal); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    pr = N.prof()
    pair = os.environ.get("FA100_PAIR") == "1"  # CTA pairs: 512 rows per work item, 74 pairs walk the list
    items = h * b * (s // (512 if pair else 256)); grid = min(items, 74 if pair else 148)
    mine = sum(1 for wi in range(items) if (wi * grid + (grid - 1 if wi & 1 else 0)) < items)
    kb = mine * (s // 128)
    w0 = pr[0]
    tot = sum(w0) - w0[3]
    print(f"b={b:3d} {ms:7.3f} ms  {b*h*2*s*s*256/ms/1e9:7.1f} TFLOP/s  CTA0 items {mine:3d}  clk/kb: total {tot/kb:7.1f}  " +
          " ".join(f"{x/kb:6.1f}" for x in w0), flush=True)
