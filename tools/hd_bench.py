import sys, torch
sys.path[:0] = ["/root/repo/flash-attention-annotated_b200"]
from flash_attn_interface import flash_attn_func
def t(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
for d, dv in ((256, 256), (192, 128), (192, 192)):
    for causal in (False, True):
        b, s, h, hk = 4, 8192, 16, 4
        q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
        k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
        v = torch.randn(b, s, hk, dv, device="cuda", dtype=torch.bfloat16)
        fl = b * h * 2 * s * s * (d + dv) / (2 if causal else 1)
        ms = t(lambda: flash_attn_func(q, k, v, causal=causal))
        line = f"d={d} dv={dv} causal={causal}: fa100 {fl/ms/1e9:.0f} TFLOP/s"
        try:
            import torch.nn.functional as F
            from torch.nn.attention import SDPBackend, sdpa_kernel
            if d == dv:
                qt, kt, vt = (x.transpose(1, 2) for x in (q, k, v))
                with sdpa_kernel(SDPBackend.CUDNN_ATTENTION):
                    ms2 = t(lambda: F.scaled_dot_product_attention(qt, kt, vt, is_causal=causal, enable_gqa=True))
                line += f"  cudnn {fl/ms2/1e9:.0f}"
        except Exception as e:
            line += f"  cudnn unavailable: {str(e)[:80]}"
        print(line, flush=True)
