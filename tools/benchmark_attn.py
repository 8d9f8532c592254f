"""The reference's headline sweep (hopper/benchmark_attn.py:229-420) for fa100: dim 2048 split into heads of 64 / 128 / 256,
(batch, seqlen) pairs with batch * seqlen = 32 k tokens, causal and non-causal, forward and backward, in TFLOP/s with the
reference's FLOP count (`flops`, :62-73; backward = 2.5 x forward).  cuDNN (torch SDPA backend) and flash_attn 2.8 are timed
beside it where they run.  python tools/benchmark_attn.py [--fp8] [--hdim 64] [--fwd-only] > profiles/<tag>_benchmark_attn_sweep.txt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch
import torch.nn.functional as F
from flash_attn_interface import flash_attn_func

FP8 = "--fp8" in sys.argv
FWD_ONLY = "--fwd-only" in sys.argv
HDIMS = (int(sys.argv[sys.argv.index("--hdim") + 1]),) if "--hdim" in sys.argv else (64, 128, 256)
dev = "cuda"


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def flops(b, h, sq, sk, d, dv, causal):
    avg = (max(0, sk - sq) + sk) / 2 if causal else sk
    return b * h * 2 * sq * avg * (d + dv)


def cudnn_fwd(q, k, v, causal):
    from torch.nn.attention import SDPBackend, sdpa_kernel
    qt, kt, vt = (x.transpose(1, 2) for x in (q, k, v))
    with sdpa_kernel(SDPBackend.CUDNN_ATTENTION):
        return timeit(lambda: F.scaled_dot_product_attention(qt, kt, vt, is_causal=causal))


def fa2_fwd(q, k, v, causal):
    sys_path = list(sys.path)
    try:
        sys.path[:] = [p for p in sys.path if "flash-attention-annotated_b200" not in p]
        import flash_attn as fa2
        return timeit(lambda: fa2.flash_attn_func(q, k, v, causal=causal))
    finally:
        sys.path[:] = sys_path


print(f"# fa100 sweep ({'e4m3' if FP8 else 'bf16'}), dim 2048, batch x seqlen = 32768 tokens; TFLOP/s (backward counted as 2.5 x forward)")
print(f"{'hdim':>4} {'causal':>6} {'batch':>5} {'seqlen':>6} | {'fa100 fwd':>9} {'fa100 bwd':>9} | {'cuDNN fwd':>9} {'FA2 fwd':>8}")
for headdim in HDIMS:
    nheads = 2048 // headdim
    for causal in (False, True):
        for batch, seqlen in ((32, 1024), (16, 2048), (8, 4096), (4, 8192), (2, 16384)):
            q = torch.randn(batch, seqlen, nheads, headdim, device=dev, dtype=torch.bfloat16)
            k, v = torch.randn_like(q), torch.randn_like(q)
            fl = flops(batch, nheads, seqlen, seqlen, headdim, headdim, causal)
            tf = lambda ms: f"{fl / ms / 1e9:9.0f}" if ms else f"{'-':>9}"  # noqa: E731
            if FP8:
                q8, k8, v8 = (x.to(torch.float8_e4m3fn) for x in (q, k, v))
                ms_f = timeit(lambda: flash_attn_func(q8, k8, v8, causal=causal))
                ms_b = None
            else:
                ms_f = timeit(lambda: flash_attn_func(q, k, v, causal=causal))
                ms_b = None
                if headdim <= 128 and not FWD_ONLY:
                    out, lse, *_ = torch.ops.flash_attn_3.fwd(q, k, v, is_causal=causal)
                    g = torch.randn_like(out)
                    ms_b = timeit(lambda: torch.ops.flash_attn_3.bwd(g, q, k, v, out, lse, None, None, None, None, None, None, None, None, None, None, causal), iters=5)
            ms_c = ms_2 = None
            if not FP8:
                try:
                    ms_c = cudnn_fwd(q, k, v, causal)
                except Exception:  # noqa: BLE001  (no cuDNN kernel for this shape on this build)
                    ms_c = None
                try:
                    ms_2 = fa2_fwd(q, k, v, causal)
                except Exception:  # noqa: BLE001
                    ms_2 = None
            bw = f"{2.5 * fl / ms_b / 1e9:9.0f}" if ms_b else f"{'-':>9}"
            print(f"{headdim:>4} {str(causal):>6} {batch:>5} {seqlen:>6} | {tf(ms_f)} {bw} | {tf(ms_c)} {tf(ms_2)[1:]}", flush=True)
            del q, k, v
            torch.cuda.empty_cache()
