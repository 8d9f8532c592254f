"""SURVEY.md 8(d) "also timed beside ours on the GPU": library attention kernels on the same shapes and the same box.
Each entry is a number or "unavailable: <reason>".  Library code only -- nothing here is on the product path.

  python tools/compare_libs.py > gpurun_out/compare.json
"""
import json
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402


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


def attempt(res, name, fn):
    try:
        res[name] = fn()
    except Exception as e:  # noqa: BLE001
        res[name] = "unavailable: " + (str(e).splitlines()[0][:160] if str(e) else type(e).__name__)
        traceback.print_exc(file=sys.stderr)


def main():
    dev = "cuda"
    out = {"device": torch.cuda.get_device_name(0)}
    # ---- C2: bf16 b8 32q/8kv s8192 d128 causal
    b, s, h, hk, d = 8, 8192, 32, 8, 128
    q = torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16)
    v = torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16)
    fl = b * h * 2 * s * (s / 2) * 2 * d
    c2 = {}

    def ours():
        from flash_attn_interface import flash_attn_func
        return fl / timeit(lambda: flash_attn_func(q, k, v, causal=True)) / 1e9

    def sdpa(backend):
        from torch.nn.attention import SDPBackend, sdpa_kernel
        qt, kt, vt = (x.transpose(1, 2) for x in (q, k, v))

        def run():
            with sdpa_kernel(backend):
                return F.scaled_dot_product_attention(qt, kt, vt, is_causal=True, enable_gqa=True)
        return fl / timeit(run) / 1e9

    def fa2():
        import flash_attn
        return fl / timeit(lambda: flash_attn.flash_attn_func(q, k, v, causal=True)) / 1e9

    def flashinfer_prefill():
        import flashinfer
        qq, kk, vv = q[0], k[0], v[0]  # (s, h, d) single sequence, scaled to the batch
        ms = timeit(lambda: flashinfer.single_prefill_with_kv_cache(qq, kk, vv, causal=True))
        return fl / b / ms / 1e9

    attempt(c2, "fa100 (this repo) TFLOP/s", ours)
    from torch.nn.attention import SDPBackend
    attempt(c2, "torch SDPA cuDNN TFLOP/s", lambda: sdpa(SDPBackend.CUDNN_ATTENTION))
    attempt(c2, "torch SDPA flash TFLOP/s", lambda: sdpa(SDPBackend.FLASH_ATTENTION))
    attempt(c2, "flash_attn 2.8 (FA2 kernels) TFLOP/s", fa2)
    attempt(c2, "flashinfer single_prefill TFLOP/s", flashinfer_prefill)
    out["C2 bf16 b8 32q/8kv s8192 d128 causal"] = c2
    del q, k, v
    torch.cuda.empty_cache()

    # ---- the reference's own Blackwell kernel (CuTe-DSL FlashAttentionForwardSm100, flash_attn/cute/interface.py), installed
    # unmodified by tools/install_ref_cute.sh; a subprocess so that its `flash_attn` package does not collide with the FA2 wheel
    import subprocess
    ref_dir = os.path.join(ROOT, "baseline", "_ref", "flash_attn", "cute")
    if not os.path.isdir(ref_dir):
        out["reference CuTe-DSL FlashAttentionForwardSm100"] = "unavailable: baseline/_ref/flash_attn/cute missing (run tools/install_ref_cute.sh)"
    else:
        try:
            r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "time_ref_cute.py")], capture_output=True, text=True, timeout=900)
            last = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
            out["reference CuTe-DSL FlashAttentionForwardSm100"] = json.loads(last[-1]) if last else "unavailable: rc=%d %s" % (r.returncode, r.stderr[-1500:])
        except Exception as e:  # noqa: BLE001
            out["reference CuTe-DSL FlashAttentionForwardSm100"] = "unavailable: " + repr(e)[:300]

    # ---- C4: decode b64 32q/8kv sq1 sk32768 d128 page 256
    b, sk, h, hk, d, ps = 64, 32768, 32, 8, 128, 256
    npages = b * sk // ps
    q = torch.randn(b, 1, h, d, device=dev, dtype=torch.bfloat16)
    kp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
    table = torch.randperm(npages, device=dev).to(torch.int32).view(b, sk // ps)
    lens = torch.full((b,), sk, dtype=torch.int32, device=dev)
    nbytes = 2 * b * sk * hk * d * 2 + 2 * b * h * d * 2 + table.numel() * 4
    c4 = {}

    def ours_dec():
        from flash_attn_interface import flash_attn_with_kvcache
        return nbytes / timeit(lambda: flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens, page_table=table)) / 1e6

    def fa2_dec():
        import flash_attn
        return nbytes / timeit(lambda: flash_attn.flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens, block_table=table)) / 1e6

    def flashinfer_dec():
        import flashinfer
        ws = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD")
        indptr = torch.arange(0, b + 1, device=dev, dtype=torch.int32) * (sk // ps)
        last = torch.full((b,), ps, dtype=torch.int32, device=dev)
        w.plan(indptr, table.flatten().contiguous(), last, h, hk, d, ps, q_data_type=torch.bfloat16)
        qq = q[:, 0].contiguous()
        return nbytes / timeit(lambda: w.run(qq, (kp, vp))) / 1e6

    attempt(c4, "fa100 (this repo) GB/s", ours_dec)
    attempt(c4, "flash_attn 2.8 flash_attn_with_kvcache GB/s", fa2_dec)
    attempt(c4, "flashinfer BatchDecodeWithPagedKVCache GB/s", flashinfer_dec)
    out["C4 decode b64 32q/8kv sk32768 d128 page256"] = c4
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
