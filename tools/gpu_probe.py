"""GPU bring-up ladder: simplest case first, never stops at a failure, prints one line per case.
Usage (on the GPU box): python tools/gpu_probe.py [substring-filter]"""
import os
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200"), os.path.join(ROOT, "oracle")]
import torch

import attention_oracle as A
import fa100_native as N
from flash_attn_interface import flash_attn_func, flash_attn_varlen_func, flash_attn_with_kvcache, flash_attn_combine

dev = "cuda"
FILT = sys.argv[1] if len(sys.argv) > 1 else ""


def report(name, out, ref, low, lse=None, lse_ref=None, rtol=2.0):
    torch.cuda.synchronize()
    hang = N.hang_info()
    err = (out.float().cpu() - ref.float()).abs().max().item()
    tol = A.fwd_tolerance(ref, low, rtol)
    msg = f"{name:42s} err={err:.3e} tol={tol:.3e} {'OK ' if err <= tol else 'BAD'}"
    if lse is not None:
        fin = torch.isfinite(lse_ref)
        le = ((lse.cpu()[fin] - lse_ref[fin]).abs() / lse_ref[fin].abs().clamp_min(1.0)).max().item() if fin.any() else 0.0
        inf_ok = torch.equal(torch.isneginf(lse.cpu()), torch.isneginf(lse_ref))
        msg += f" lse={le:.2e} {'OK ' if le <= 1e-3 and inf_ok else 'BAD'}"
    if hang[0]:
        msg += f" HANG tag={hang[0] & 0xffff:#x} blk={hang[1]} thr={hang[2]} par={hang[3]}"
    if not torch.isfinite(out.float()).all():
        msg += " NONFINITE"
    print(msg, flush=True)


def dense(name, b, sq, sk, h, hk, d, dtype=torch.bfloat16, seed=0, **kw):
    if FILT not in name:
        return
    try:
        g = torch.Generator().manual_seed(seed)
        q = torch.randn(b, sq, h, d, generator=g).to(dtype)
        k = torch.randn(b, sk, hk, d, generator=g).to(dtype)
        v = torch.randn(b, sk, hk, d, generator=g).to(dtype)
        okw = {x: kw[x] for x in ("causal", "window_size", "attention_chunk", "softcap") if x in kw}
        ref, lse_ref = A.attention_oracle(q, k, v, **okw)
        low, _ = A.attention_oracle(q, k, v, upcast=False, reorder_ops=True, **okw)
        out, lse = flash_attn_func(q.to(dev), k.to(dev), v.to(dev), return_attn_probs=True, **kw)
        report(name, out, ref, low, lse, lse_ref, rtol=3.0 if kw.get("softcap", 0) else 2.0)
    except Exception as e:  # noqa: BLE001
        print(f"{name:42s} EXC {type(e).__name__}: {e}", flush=True)
        traceback.print_exc()


dense("a1 d128 1x128x128", 1, 128, 128, 1, 1, 128)
dense("a2 d128 1x256x256", 1, 256, 256, 1, 1, 128)
dense("a3 d128 1x128x512", 1, 128, 512, 1, 1, 128)
dense("a4 d128 causal 512", 1, 512, 512, 2, 2, 128, causal=True)
dense("a5 d128 odd 113x203 gqa", 2, 113, 203, 6, 2, 128)
dense("a6 d128 odd causal 113x203 gqa", 2, 113, 203, 6, 2, 128, causal=True)
dense("a7 d128 causal 1023x1024", 2, 1023, 1024, 4, 1, 128, causal=True)
dense("a8 d128 causal sq>sk 600x300", 1, 600, 300, 2, 2, 128, causal=True)
dense("b1 d64 256", 1, 256, 256, 2, 2, 64)
dense("b2 d64 causal odd", 2, 239, 400, 4, 2, 64, causal=True)
dense("b3 d96 causal", 1, 300, 300, 2, 1, 96, causal=True)
dense("c1 fp16 d128 causal", 1, 384, 384, 2, 2, 128, dtype=torch.float16, causal=True)
dense("d1 local (100,50)", 1, 500, 700, 2, 1, 128, window_size=(100, 50))
dense("d2 local (300,0)", 1, 800, 800, 2, 1, 128, window_size=(300, 0))
dense("d3 chunk 128", 1, 500, 500, 2, 1, 128, attention_chunk=128)
dense("d4 softcap", 1, 300, 300, 2, 1, 128, softcap=15.0, causal=True)
dense("e1 split3", 1, 200, 1500, 2, 1, 128, num_splits=3)
dense("e2 split3 causal", 1, 700, 1500, 2, 1, 128, num_splits=3, causal=True)
dense("e3 packgqa", 1, 70, 600, 8, 2, 128, pack_gqa=True, causal=True)
dense("e4 packgqa G3 split", 2, 33, 900, 6, 2, 128, pack_gqa=True, num_splits=2)
dense("e5 decode sq1 packed split", 3, 1, 3000, 8, 2, 128, pack_gqa=True, num_splits=4)
dense("f1 big causal 2048", 2, 2048, 2048, 8, 2, 128, causal=True)


def varlen(name, lens_q, lens_k, h, hk, d, **kw):
    if FILT not in name:
        return
    try:
        g = torch.Generator().manual_seed(1)
        b = len(lens_q)
        cq = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32)
        ck = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32)
        q = torch.randn(int(cq[-1]), h, d, generator=g).to(torch.bfloat16)
        k = torch.randn(int(ck[-1]), hk, d, generator=g).to(torch.bfloat16)
        v = torch.randn(int(ck[-1]), hk, d, generator=g).to(torch.bfloat16)
        out, lse = flash_attn_varlen_func(q.to(dev), k.to(dev), v.to(dev), cq.to(dev), ck.to(dev), max(lens_q), max(lens_k),
                                          return_attn_probs=True, **kw)
        refs, lows, lses = [], [], []
        for i in range(b):
            qi, ki, vi = q[cq[i]:cq[i + 1]][None], k[ck[i]:ck[i + 1]][None], v[ck[i]:ck[i + 1]][None]
            if lens_k[i] == 0:
                refs.append(torch.zeros_like(qi[0])); lows.append(torch.zeros_like(qi[0])); lses.append(torch.full((h, lens_q[i]), float("-inf")))
                continue
            r, l = A.attention_oracle(qi, ki, vi, causal=kw.get("causal", False))
            lo, _ = A.attention_oracle(qi, ki, vi, causal=kw.get("causal", False), upcast=False, reorder_ops=True)
            refs.append(r[0]); lows.append(lo[0]); lses.append(l[0])
        report(name, out, torch.cat(refs), torch.cat(lows), lse, torch.cat(lses, dim=1))
    except Exception as e:  # noqa: BLE001
        print(f"{name:42s} EXC {type(e).__name__}: {e}", flush=True)
        traceback.print_exc()


varlen("g1 varlen", [100, 257, 31], [100, 300, 77], 4, 2, 128)
varlen("g2 varlen causal zero-len", [128, 5, 300], [128, 0, 511], 4, 1, 128, causal=True)


def kvcache(name, b, sq, sk_cache, h, hk, d, page_size=None, causal=True, **kw):
    if FILT not in name:
        return
    try:
        g = torch.Generator().manual_seed(2)
        q = torch.randn(b, sq, h, d, generator=g).to(torch.bfloat16)
        seqlens = torch.randint(sk_cache // 2, sk_cache + 1, (b,), generator=g, dtype=torch.int32)
        if page_size is None:
            kc = torch.randn(b, sk_cache, hk, d, generator=g).to(torch.bfloat16)
            vc = torch.randn(b, sk_cache, hk, d, generator=g).to(torch.bfloat16)
            kd, vd, table = kc, vc, None
            kin, vin = kc, vc
        else:
            kd, vd, table, kin, vin = A.paged_kvcache(sk_cache, page_size, b, hk, d, d, torch.bfloat16, generator=g)
        out, lse, *_ = flash_attn_with_kvcache(q.to(dev), kin.to(dev), vin.to(dev), cache_seqlens=seqlens.to(dev),
                                               page_table=None if table is None else table.to(dev), causal=causal,
                                               return_softmax_lse=True, **kw)
        ref, lse_ref = A.attention_oracle(q, kd, vd, k_lens=seqlens, causal=causal)
        low, _ = A.attention_oracle(q, kd, vd, k_lens=seqlens, causal=causal, upcast=False, reorder_ops=True)
        report(name, out, ref, low, lse, lse_ref)
    except Exception as e:  # noqa: BLE001
        print(f"{name:42s} EXC {type(e).__name__}: {e}", flush=True)
        traceback.print_exc()


kvcache("h1 kvcache dense sq1", 4, 1, 2000, 8, 2, 128)
kvcache("h2 kvcache dense sq16 nosplit", 2, 16, 1500, 4, 4, 128, num_splits=1)
kvcache("h3 kvcache paged256 sq1", 4, 1, 2000, 8, 2, 128, page_size=256)
kvcache("h4 kvcache paged128 sq64", 2, 64, 1000, 4, 1, 128, page_size=128, num_splits=1)
kvcache("h5 kvcache paged64 sq1", 2, 1, 700, 8, 2, 128, page_size=64)
kvcache("h6 kvcache paged16 sq3", 2, 3, 300, 4, 2, 128, page_size=16)


def fp8(name, b, sq, sk, h, hk, d, causal):
    if FILT not in name:
        return
    try:
        g = torch.Generator().manual_seed(3)
        q = torch.randn(b, sq, h, d, generator=g).to(torch.float8_e4m3fn)
        k = torch.randn(b, sk, hk, d, generator=g).to(torch.float8_e4m3fn)
        v = torch.randn(b, sk, hk, d, generator=g).to(torch.float8_e4m3fn)
        ds = [torch.full((b, hk), x) for x in (0.5, 0.75, 1.25)]
        qr, kr, vr = (x.to(torch.bfloat16) for x in (q, k, v))
        ref, lse_ref = A.attention_oracle(qr, kr, vr, causal=causal, q_descale=ds[0], k_descale=ds[1], v_descale=ds[2])
        low, _ = A.attention_oracle(qr, kr, vr, causal=causal, q_descale=ds[0], k_descale=ds[1], v_descale=ds[2], upcast=False,
                                    reorder_ops=True, intermediate_dtype=torch.float8_e4m3fn)
        out, lse = flash_attn_func(q.to(dev), k.to(dev), v.to(dev), causal=causal, q_descale=ds[0].to(dev), k_descale=ds[1].to(dev),
                                   v_descale=ds[2].to(dev), return_attn_probs=True)
        report(name, out, ref, low, lse, lse_ref, rtol=4.0)
    except Exception as e:  # noqa: BLE001
        print(f"{name:42s} EXC {type(e).__name__}: {e}", flush=True)
        traceback.print_exc()


fp8("i1 fp8 128", 1, 128, 128, 1, 1, 128, False)
fp8("i2 fp8 causal 700", 2, 700, 700, 4, 2, 128, True)

if FILT in "z timing":
    try:
        for (b, s, h, hk, causal) in ((8, 8192, 32, 8, True), (4, 8192, 32, 8, False)):
            q = torch.randn(b, s, h, 128, device=dev, dtype=torch.bfloat16)
            k = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16)
            v = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16)
            for _ in range(3):
                flash_attn_func(q, k, v, causal=causal)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                flash_attn_func(q, k, v, causal=causal)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            fl = A.flops(b, h, s, s, 128, 128, causal)
            print(f"z timing b{b} s{s} h{h}/{hk} causal={causal}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOPS  hang={N.hang_info()}", flush=True)
    except Exception as e:  # noqa: BLE001
        print("timing EXC", e, flush=True)
        traceback.print_exc()
