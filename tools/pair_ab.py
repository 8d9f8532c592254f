"""1-CTA against 2-CTA (cta_group::2) kernel in one process, interleaved bursts on the same tensors (FA100_PAIR is read per call).
  python tools/pair_ab.py"""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
from flash_attn_interface import flash_attn_func
dev = "cuda"
def burst(fn, n=10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for name, (b, s, h, hk, causal) in (("C2 causal b8 32/8 s8192", (8, 8192, 32, 8, True)), ("non-causal b4 32/8 s8192", (4, 8192, 32, 8, False)),
                                     ("non-causal b2 16/16 s16384", (2, 16384, 16, 16, False)), ("causal b1 32/32 s32768", (1, 32768, 32, 32, True))):
    q = torch.randn(b, s, h, 128, device=dev, dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16)
    v = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16)
    fn = lambda: flash_attn_func(q, k, v, causal=causal)
    res = {"0": [], "1": []}
    for rep in range(6):
        for mode in ("0", "1") if rep % 2 == 0 else ("1", "0"):
            os.environ["FA100_PAIR"] = mode
            fn(); fn()
            res[mode].append(burst(fn))
    fl = b * h * 2 * s * (s / 2 if causal else s) * 256
    tf = {m: [fl / x / 1e9 for x in v_] for m, v_ in res.items()}
    print(f"{name:30s} 1-CTA median {statistics.median(tf['0']):7.1f} (min {min(tf['0']):7.1f} max {max(tf['0']):7.1f})   2-CTA median {statistics.median(tf['1']):7.1f} (min {min(tf['1']):7.1f} max {max(tf['1']):7.1f}) TFLOP/s", flush=True)
