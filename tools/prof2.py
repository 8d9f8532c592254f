"""Cycle buckets of the v2 prefill kernel (cycle-bucket build, CTA 0): per role, clk per key block of one Q tile.
  FA100_LIB=libfa100_prof.so python tools/prof2.py [--fp8] [--causal] [--legacy]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
os.environ.setdefault("FA100_LIB", "libfa100_prof.so")
if "--legacy" not in sys.argv:
    os.environ["FA100_FOUR_ROLE"] = "1"
import torch
import fa100_native as N
from flash_attn_interface import flash_attn_func
dt = torch.float8_e4m3fn if "--fp8" in sys.argv else torch.bfloat16
b, s, h, hk = 8, 8192, 16, 4
causal = "--causal" in sys.argv
q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16).to(dt)
k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16).to(dt)
v = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16).to(dt)
flash_attn_func(q, k, v, causal=causal); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); flash_attn_func(q, k, v, causal=causal); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
pr = N.prof(16 * 8)
items = h * b * (s // 256); grid = min(items, 148)
mine = sum(1 for wi in range(items) if (wi * grid + (grid - 1 if wi & 1 else 0)) < items)
kb = mine * (s // 128) * (0.5 if causal else 1.0)
fl = b * h * 2 * s * (s / 2 if causal else s) * 256
print(f"{'fp8' if dt != torch.bfloat16 else 'bf16'} causal={causal} legacy={'--legacy' in sys.argv}: {ms:7.3f} ms {fl/ms/1e9:7.1f} TFLOP/s; CTA0 {mine} items, ~{kb:.0f} key blocks per tile")
names = {0: "softmax t0 w0", 4: "softmax t1 w4", 8: "corr/QK w8", 9: "w9", 10: "w10", 12: "mma w12", 13: "producer w13"}
for w, nm in names.items():
    if w < len(pr) and sum(pr[w]):
        print(f"  {nm:14s} total {sum(pr[w])/kb:7.1f} | " + " ".join(f"{i}:{x/kb:6.1f}" for i, x in enumerate(pr[w])))
