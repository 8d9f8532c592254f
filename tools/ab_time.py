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
D = int(os.environ.get("AB_HEADDIM", "128"))  # AB_HEADDIM=64: the sweep's hdim-64 shapes (MHA, 32 heads) instead of C2 / C3
shapes = ((8, 8192, 32, 8, True, torch.bfloat16), (4, 8192, 32, 8, False, torch.bfloat16), (4, 16384, 16, 16, True, torch.float8_e4m3fn), (4, 16384, 16, 16, False, torch.float8_e4m3fn))
if D != 128: shapes = ((4, 8192, 32, 32, True, torch.bfloat16), (4, 8192, 32, 32, False, torch.bfloat16), (16, 2048, 32, 32, True, torch.bfloat16), (2, 16384, 32, 32, False, torch.bfloat16))
for (b, s, h, hk, causal, dt) in shapes:
    q = torch.randn(b, s, h, D, device=dev, dtype=torch.bfloat16).to(dt)
    k = torch.randn(b, s, hk, D, device=dev, dtype=torch.bfloat16).to(dt)
    v = torch.randn(b, s, hk, D, device=dev, dtype=torch.bfloat16).to(dt)
    for _ in range(3): flash_attn_func(q, k, v, causal=causal)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): flash_attn_func(q, k, v, causal=causal)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    fl = b * h * 2 * s * (s / 2 if causal else s) * 2 * D
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
heads = ("bf16 causal", "bf16 non-causal", "fp8 causal", "fp8 non-causal") if os.environ.get("AB_HEADDIM", "128") == "128" else ("8k causal", "8k non-causal", "2k causal", "16k non-causal")
print(f"{'library':28s} " + " ".join(f"{x:>22s}" for x in heads) + f"   (TFLOPS: best / median of {rounds})")
for l in libs:
    if out[l]:
        cols = list(zip(*out[l]))
        print(f"{l:28s} " + " ".join(f"{max(c):10.1f} /{statistics.median(c):10.1f}" for c in cols))
