"""Times the reference's own Blackwell kernel -- CuTe-DSL FlashAttentionForwardSm100 through its stock entry
flash_attn/cute/interface.py (flash_attn_func / flash_attn_varlen_func) -- installed UNMODIFIED by
tools/install_ref_cute.sh into baseline/_ref.  Run as a subprocess of tools/compare_libs.py (its own sys.path, so the
FA2 wheel's `flash_attn` package does not shadow it).  Prints one JSON object; on failure the JIT traceback's tail.
Library-side comparison only: nothing here is on the product path."""
import json
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
import torch  # noqa: E402


def timeit(fn, warmup=3, iters=10):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    res = {}
    try:
        import cutlass
        res["cutlass_dsl"] = getattr(cutlass, "__version__", "?")
        from flash_attn.cute.interface import flash_attn_func, flash_attn_varlen_func
    except Exception:  # noqa: BLE001
        res["import"] = "unavailable: " + traceback.format_exc()[-1200:]
        print(json.dumps(res))
        return
    dev = "cuda"
    torch.manual_seed(0)
    # C2: bf16 b8 32q/8kv s8192 d128 causal
    b, s, h, hk, d = 8, 8192, 32, 8, 128
    q = torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16)
    v = torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16)
    fl = b * h * 2 * s * (s / 2) * 2 * d
    try:
        out = flash_attn_func(q, k, v, causal=True)
        out = out[0] if isinstance(out, tuple) else out
        torch.cuda.synchronize()
        res["C2 TFLOP/s"] = fl / timeit(lambda: flash_attn_func(q, k, v, causal=True)) / 1e9
        # cross-check against this repo's kernel on the same inputs (max abs difference of two bf16 outputs)
        sys.path.insert(0, os.path.join(ROOT, "flash-attention-annotated_b200"))
        from flash_attn_interface import flash_attn_func as ours
        res["C2 max|ref_cute - fa100|"] = (out.float() - ours(q, k, v, causal=True).float()).abs().max().item()
    except Exception:  # noqa: BLE001
        res["C2 TFLOP/s"] = "unavailable: " + traceback.format_exc()[-1500:]
    # MHA non-causal (no mask: the exp2-emulation path of the reference is on)
    try:
        qm = torch.randn(4, 8192, 16, d, device=dev, dtype=torch.bfloat16)
        km, vm = torch.randn_like(qm), torch.randn_like(qm)
        flm = 4 * 16 * 2 * 8192 * 8192 * 2 * d
        flash_attn_func(qm, km, vm, causal=False)
        res["MHA b4 h16 s8192 non-causal TFLOP/s"] = flm / timeit(lambda: flash_attn_func(qm, km, vm, causal=False)) / 1e9
    except Exception:  # noqa: BLE001
        res["MHA b4 h16 s8192 non-causal TFLOP/s"] = "unavailable: " + traceback.format_exc()[-1500:]
    # varlen: 8 sequences of 1024..8192 tokens packed, causal, same heads as C2
    try:
        lens = torch.tensor([8192, 1024, 4096, 2048, 6144, 3072, 7168, 5120], dtype=torch.int32)
        cu = torch.zeros(9, dtype=torch.int32)
        cu[1:] = lens.cumsum(0)
        cu = cu.to(dev)
        tot = int(lens.sum())
        qv = torch.randn(tot, h, d, device=dev, dtype=torch.bfloat16)
        kv = torch.randn(tot, hk, d, device=dev, dtype=torch.bfloat16)
        vv = torch.randn(tot, hk, d, device=dev, dtype=torch.bfloat16)
        flv = float(sum(h * 2 * int(n) * (int(n) / 2) * 2 * d for n in lens))
        run = lambda: flash_attn_varlen_func(qv, kv, vv, cu_seqlens_q=cu, cu_seqlens_k=cu, causal=True)  # noqa: E731
        run()
        res["varlen 8 seqs 1k..8k causal TFLOP/s"] = flv / timeit(run) / 1e9
    except Exception:  # noqa: BLE001
        res["varlen 8 seqs 1k..8k causal TFLOP/s"] = "unavailable: " + traceback.format_exc()[-1500:]
    print(json.dumps(res))


if __name__ == "__main__":
    main()
