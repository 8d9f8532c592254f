"""Cycle buckets of the default forward kernel on a short-sequence shape (cycle-bucket build, CTA 0).
  FA100_LIB=libfa100_prof.so python tools/prof_short.py B S H D causal(0/1)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
os.environ.setdefault("FA100_LIB", "libfa100_prof.so")
import torch
import fa100_native as N
from flash_attn_interface import flash_attn_func
b, s, h, d, causal = (int(x) for x in sys.argv[1:6])
q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16); k = torch.randn_like(q); v = torch.randn_like(q)
flash_attn_func(q, k, v, causal=bool(causal)); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); flash_attn_func(q, k, v, causal=bool(causal)); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
pr = N.prof(20 * 8)
items = b * h * ((s + 255) // 256); grid = min(items, 148)
mine = len([r for r in range((items + grid - 1) // grid + 1) if (r * grid + (grid - 1 if r & 1 else 0)) < items])
print(f"b{b} s{s} h{h} d{d} causal={causal}: {ms*1e3:.1f} us; CTA0 ~{mine} items; clk per ITEM per bucket")
print("  buckets: 0 s_full wait, 1 S load, 2 s_free+mask+max, 3 rescale, 4 exp, 5 st issue, 6 st wait+publish, 7 loop top / geometry / epilogue")
for w in (0, 4, 8, 9, 10, 12, 13):
    if w < len(pr) and sum(pr[w]):
        print(f"  warp {w:2d} total {sum(pr[w])/mine:8.0f} | " + " ".join(f"{i}:{x/mine:7.0f}" for i, x in enumerate(pr[w])))
