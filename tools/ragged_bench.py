"""Ragged decode / ragged varlen prefill with and without the device-built schedule (FA100_NO_SCHED=1 = static schedule).
  python tools/ragged_bench.py"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, torch
sys.path[:0] = [os.path.join(%r, "flash-attention-annotated_b200")]
from flash_attn_interface import flash_attn_with_kvcache, flash_attn_varlen_func
dev = "cuda"; g = torch.Generator(device=dev).manual_seed(0)
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
# C4 ragged variant: b64 32q/8kv sq1, sk in [16384, 32768], page 256
b, h, hk, d, ps, cap = 64, 32, 8, 128, 256, 32768
npages = b * cap // ps
kp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
vp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
table = torch.randperm(npages, device=dev).to(torch.int32).view(b, cap // ps)
lens = torch.randint(16384, 32769, (b,), device=dev, generator=g).to(torch.int32)
q = torch.randn(b, 1, h, d, device=dev, dtype=torch.bfloat16)
ms = timeit(lambda: flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens, page_table=table))
byt = int(lens.sum()) * hk * 2 * d * 2
print(f"RES ragged decode {ms:.4f} ms {byt / ms / 1e6:.1f} GB/s")
# very ragged decode: a few long sequences among short ones
lens2 = torch.full((b,), 512, device=dev, dtype=torch.int32); lens2[::8] = 32768
ms = timeit(lambda: flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens2, page_table=table))
byt = int(lens2.sum()) * hk * 2 * d * 2
print(f"RES skewed decode {ms:.4f} ms {byt / ms / 1e6:.1f} GB/s")
# ragged varlen prefill: 32 sequences, 256..8192 tokens, causal
lq = torch.randint(256, 8193, (32,), device=dev, generator=g).to(torch.int32)
cu = torch.nn.functional.pad(lq.cumsum(0), (1, 0)).to(torch.int32)
tot = int(cu[-1])
qv = torch.randn(tot, h, d, device=dev, dtype=torch.bfloat16); kv = torch.randn(tot, hk, d, device=dev, dtype=torch.bfloat16); vv = torch.randn(tot, hk, d, device=dev, dtype=torch.bfloat16)
ms = timeit(lambda: flash_attn_varlen_func(qv, kv, vv, cu, cu, 8192, 8192, causal=True))
fl = float((lq.double() ** 2).sum()) * h * 2 * d * 2 / 2
print(f"RES ragged varlen prefill {ms:.4f} ms {fl / ms / 1e9:.1f} TFLOP/s")
''' % ROOT
for env in ({}, {"FA100_NO_SCHED": "1"}):
    p = subprocess.run([sys.executable, "-c", CHILD], env=dict(os.environ, **env), capture_output=True, text=True, timeout=300)
    print("static schedule" if env else "device-built schedule")
    print("\n".join(l for l in p.stdout.splitlines() if l.startswith("RES")) or p.stderr[-800:])
