"""A/B timing of library builds on the same box: python tools/ab_time.py libfa100_a.so libfa100_b.so ...
Each library is timed in its own subprocess (FA100_LIB), rounds interleaved, best and median of the rounds reported."""
import os, subprocess, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, torch
sys.path[:0] = [os.path.join(%r, "flash-attention-annotated_b200")]
from flash_attn_interface import flash_attn_func
dev = "cuda"
res = []
for (b, s, h, hk, causal, dt) in ((8, 8192, 32, 8, True, torch.bfloat16), (4, 8192, 32, 8, False, torch.bfloat16), (4, 16384, 16, 16, True, torch.float8_e4m3fn), (4, 16384, 16, 16, False, torch.float8_e4m3fn)):
    q = torch.randn(b, s, h, 128, device=dev, dtype=torch.bfloat16).to(dt)
    k = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16).to(dt)
    v = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16).to(dt)
    for _ in range(3): flash_attn_func(q, k, v, causal=causal)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): flash_attn_func(q, k, v, causal=causal)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    fl = b * h * 2 * s * (s / 2 if causal else s) * 256
    res.append(fl / ms / 1e9)
print("RES", *[f"{x:.1f}" for x in res])
''' % ROOT
libs = sys.argv[1:]
rounds = int(os.environ.get("AB_ROUNDS", "3"))
out = {l: [] for l in libs}
for r in range(rounds):
    for l in libs:
        lib, *envs = l.split(":")  # "libfa100.so:FA100_NO_TMA_STORE=1" = that library with an environment switch set
        p = subprocess.run([sys.executable, "-c", CHILD], env=dict(os.environ, FA100_LIB=lib, **dict(e.split("=") for e in envs)), capture_output=True, text=True, timeout=200)
        line = [x for x in p.stdout.splitlines() if x.startswith("RES")]
        if not line:
            print(l, "FAILED", p.stderr[-500:]); continue
        out[l].append([float(x) for x in line[0].split()[1:]])
print(f"{'library':28s} {'bf16 causal':>22s} {'bf16 non-causal':>22s} {'fp8 causal':>22s} {'fp8 non-causal':>22s}   (TFLOPS: best / median of {rounds})")
for l in libs:
    if out[l]:
        cols = list(zip(*out[l]))
        print(f"{l:28s} " + " ".join(f"{max(c):10.1f} /{statistics.median(c):10.1f}" for c in cols))
