"""A/B of library builds on the decode path (C4 and a b=1 decode), same box, interleaved rounds, each library in its own
subprocess (FA100_LIB):  python tools/ab_decode.py libfa100.so libfa100_nopdl.so"""
import os, subprocess, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, torch
sys.path[:0] = [os.path.join(%r, "flash-attention-annotated_b200")]
from flash_attn_interface import flash_attn_with_kvcache
dev = "cuda"
res = []
for (b, sk) in ((64, 32768), (1, 4096), (8, 8192)):
    h, hk, d, ps = 32, 8, 128, 256
    npages = b * sk // ps
    q = torch.randn(b, 1, h, d, device=dev, dtype=torch.bfloat16)
    kp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
    table = torch.randperm(npages, device=dev).to(torch.int32).view(b, sk // ps)
    lens = torch.full((b,), sk, dtype=torch.int32, device=dev)
    fn = lambda: flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens, page_table=table)
    g = torch.cuda.CUDAGraph()
    fn(); torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(10): fn()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): g.replay()
    e1.record(); torch.cuda.synchronize()
    res.append(e0.elapsed_time(e1) / 30 * 1e3)   # us per call, GPU time only (graph replay: no host gaps)
print("RES", *[f"{x:.2f}" for x in res])
''' % ROOT
libs = sys.argv[1:]
rounds = int(os.environ.get("AB_ROUNDS", "3"))
out = {l: [] for l in libs}
for r in range(rounds):
    for l in libs:
        p = subprocess.run([sys.executable, "-c", CHILD], env=dict(os.environ, FA100_LIB=l), capture_output=True, text=True, timeout=300)
        line = [x for x in p.stdout.splitlines() if x.startswith("RES")]
        if not line:
            print(l, "FAILED", p.stderr[-600:]); continue
        out[l].append([float(x) for x in line[0].split()[1:]])
print(f"{'library':24s} {'C4 b64 sk32768':>20s} {'b1 sk4096':>20s} {'b8 sk8192':>20s}   (us per call in a 10-call CUDA graph: best / median of {rounds})")
for l in libs:
    if out[l]:
        cols = list(zip(*out[l]))
        print(f"{l:24s} " + " ".join(f"{min(c):9.2f} /{statistics.median(c):9.2f}" for c in cols))
