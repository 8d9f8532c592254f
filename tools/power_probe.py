"""Power / clock behaviour under a sustained attention load (C2 shape): NVML samples every 10 ms while `which` runs for ~2 s.
  python tools/power_probe.py fa100|cudnn"""
import os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch, pynvml as nv
import torch.nn.functional as F
which = sys.argv[1] if len(sys.argv) > 1 else "fa100"
b, s, h, hk, d = 8, 8192, 32, 8, 128
q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
if which == "fa100":
    from flash_attn_interface import flash_attn_func
    run = lambda: flash_attn_func(q, k, v, causal=True)
else:
    from torch.nn.attention import SDPBackend, sdpa_kernel
    qt, kt, vt = (x.transpose(1, 2) for x in (q, k, v))
    def run():
        with sdpa_kernel(SDPBackend.CUDNN_ATTENTION):
            return F.scaled_dot_product_attention(qt, kt, vt, is_causal=True, enable_gqa=True)
nv.nvmlInit(); hnd = nv.nvmlDeviceGetHandleByIndex(0)
rows, stop = [], False
def pump():
    while not stop:
        rows.append((time.perf_counter(), nv.nvmlDeviceGetClockInfo(hnd, nv.NVML_CLOCK_SM), nv.nvmlDeviceGetPowerUsage(hnd) / 1000.0,
                     nv.nvmlDeviceGetCurrentClocksEventReasons(hnd)))
        time.sleep(0.01)
for _ in range(3): run()
torch.cuda.synchronize()
th = threading.Thread(target=pump, daemon=True); th.start()
n = 500
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n): run()
e1.record(); torch.cuda.synchronize()
stop = True; th.join()
ms = e0.elapsed_time(e1) / n
fl = b * h * 2 * s * (s / 2) * 2 * d
clk = sorted(r[1] for r in rows); pw = sorted(r[2] for r in rows)
reasons = 0
for r in rows: reasons |= r[3]
lim = nv.nvmlDeviceGetEnforcedPowerLimit(hnd) / 1000.0
print(f"{which}: {n} launches, {ms:.3f} ms each, {fl / ms / 1e9:.1f} TFLOP/s; SM clock median {clk[len(clk)//2]} MHz (min {clk[0]}, max {clk[-1]}); "
      f"power median {pw[len(pw)//2]:.0f} W (max {pw[-1]:.0f} W, limit {lim:.0f} W); reasons bitmask {reasons:#x}; samples {len(rows)}")
