"""Timeline of one CTA of the forward kernel (needs FA100_LIB=libfa100_trace.so).  Prints per-role wait/busy
statistics and the first events of a steady-state window.  Usage: FA100_LIB=libfa100_trace.so python tools/trace_timeline.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
os.environ.setdefault("FA100_LIB", "libfa100_trace.so")
import torch
import fa100_native as N
from flash_attn_interface import flash_attn_func

def _arg(name, default):
    return int(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default


b, s, h, hk, d = _arg("--b", 1), _arg("--s", 8192), _arg("--h", 16), _arg("--hk", 4), 128
causal = "--noncausal" not in sys.argv
dt = torch.float8_e4m3fn if "--fp8" in sys.argv else torch.bfloat16
q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16).to(dt)
k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16).to(dt)
v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16).to(dt)
flash_attn_func(q, k, v, causal=causal); torch.cuda.synchronize(); N.trace()
flash_attn_func(q, k, v, causal=causal); torch.cuda.synchronize()
ev = N.trace()
print("events", len(ev), "hang", N.hang_info())
pr = N.prof()
if pr:
    print("cycle buckets per warp (softmax: 0 s_full wait, 1 S load, 2 s_free+mask+max+rescale, 3 LAT QK issued->S seen, 4 exp chunks 0-1, 5 exp chunks 2-3, 6 st wait+publish, 7 rest;"
          " mma: 0 kv_full, 1 s_free, 2 p_full/o_empty, 3 issue, 4 rest, 5 LAT s_free arrive->QK start, 6 LAT p_full arrive->PV start)")
    for w_, b_ in enumerate(pr):
        tot = sum(b_) or 1
        print(f"  warp {w_:2d}: " + " ".join(f"{x:9d}({100 * x / tot:4.1f}%)" for x in b_))
NAMES = {0x10: "prod wait q_empty0", 0x11: "prod wait q_empty1", 0x20: "prod wait kv_empty(K)", 0x21: "prod wait kv_empty(V)",
         0x30: "mma wait q_full0", 0x31: "mma wait q_full1", 0x40: "mma wait kv_full(first K)", 0x41: "mma wait kv_full(V)",
         0x42: "mma wait kv_full(K next)", 0x48: "mma wait s_free (QK0)", 0x49: "mma wait s_free (QK1)", 0x78: "smx wait pv_done0", 0x79: "smx wait pv_done1", 0x50: "mma wait o_empty0", 0x51: "mma wait o_empty1", 0x60: "mma wait p_full0",
         0x61: "mma wait p_full1", 0x70: "smx wait s_full0", 0x71: "smx wait s_full1", 0x80: "smx wait o_full0",
         0x81: "smx wait o_full1", 0x90: "smx max done 0", 0x91: "smx max done 1", 0xA0: "smx P published 0", 0xA1: "smx P published 1",
         0xB0: "mma begin PV0", 0xB1: "mma begin PV1", 0xB2: "mma issued PV0", 0xB3: "mma issued PV1", 0xB4: "mma issued QK0",
         0xB5: "mma issued QK1", 0xB6: "mma committed kv_empty", 0xC0: "smx exp begin", 0xC1: "smx chunk0 stored", 0xC2: "smx chunk1 stored",
         0xC3: "smx st_wait done", 0xC4: "smx S in regs", 0xC5: "smx local max done"}
WARPS = tuple(int(x) for x in os.environ.get("TRACE_WARPS", "0,4,8").split(","))
t0 = min(e[2] for e in ev)
byw = {}
for tag, w, tb, te in ev:
    byw.setdefault(w, []).append((tag, (tb - t0) & 0xFFFFFFFF, (te - t0) & 0xFFFFFFFF))
for w in sorted(byw):
    evs = byw[w]
    span = evs[-1][2] - evs[0][1]
    agg = {}
    for tag, tb, te in evs:
        a = agg.setdefault(tag, [0, 0]); a[0] += 1; a[1] += te - tb
    print(f"warp {w}: {len(evs)} events over {span} clk")
    for tag, (n, tot) in sorted(agg.items()):
        print(f"    {NAMES.get(tag, hex(tag)):28s} n={n:5d} total wait={tot:9d} clk ({100 * tot / max(span, 1):5.1f}%) avg={tot / n:8.1f}")
# steady-state window: merge warps 0, 4, 8 (first softmax warp of each tile + MMA thread)
sel = sorted(((tb, te, tag, w) for w in WARPS for tag, tb, te in byw.get(w, [])), key=lambda x: x[1])
start = len(sel) // 2
print("\n--- steady-state window (sorted by end time): end_clk  role  event  [wait clk]")
base = sel[start][1]
for tb, te, tag, w in sel[start:start + 90]:
    print(f"{te - base:8d}  w{w:<2d} {NAMES.get(tag, hex(tag)):28s} waited {te - tb:6d}")
