"""Generate golden vectors from the REFERENCE's own oracle.  Run in the build container only:

    python tests/golden/make_golden.py

It imports ``attention_ref`` from /root/reference/hopper/test_util.py (pure PyTorch) and the
combine reference from hopper/test_flash_attn.py:1105-1114 (re-evaluated here from its three
lines because that module imports the CUDA extension at import time), evaluates them on
seeded inputs and stores inputs + outputs in ``tests/golden/attention_ref.pt``.  The GPU box
never sees /root/reference; tests read only the stored file.
"""
import os
import sys

import torch

REF = "/root/reference/hopper"
sys.path.insert(0, REF)
from test_util import attention_ref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = [
    # name, (b, sq, sk, h, hk, d, dv), dtype, kwargs
    ("plain_c1like", (2, 64, 64, 4, 4, 64, 64), torch.float32, {}),
    ("causal_gqa", (2, 48, 80, 6, 2, 32, 32), torch.bfloat16, dict(causal=True)),
    ("causal_sq_gt_sk", (1, 40, 24, 2, 1, 32, 32), torch.bfloat16, dict(causal=True)),
    ("local", (2, 33, 57, 4, 2, 32, 32), torch.float16, dict(window_size=(9, 4))),
    ("local_left_only", (1, 33, 57, 2, 2, 32, 32), torch.float16, dict(window_size=(11, 0))),
    ("chunk", (1, 50, 50, 2, 1, 32, 32), torch.bfloat16, dict(attention_chunk=16)),
    ("chunk_local", (1, 50, 70, 2, 1, 32, 32), torch.bfloat16, dict(attention_chunk=16, window_size=(20, 3))),
    ("softcap", (1, 31, 45, 2, 2, 64, 64), torch.bfloat16, dict(softcap=15.0)),
    ("softcap_causal", (1, 31, 45, 2, 2, 64, 64), torch.bfloat16, dict(softcap=15.0, causal=True)),
    ("dv_ne_d", (1, 20, 36, 2, 2, 64, 32), torch.bfloat16, dict(causal=True)),
    ("lowprec_arm", (2, 48, 80, 6, 2, 32, 32), torch.bfloat16, dict(causal=True, upcast=False, reorder_ops=True)),
    ("padmask", (3, 40, 56, 4, 2, 32, 32), torch.bfloat16, dict(causal=True, _pad=True)),
    ("padmask_local", (3, 40, 56, 4, 2, 32, 32), torch.bfloat16, dict(window_size=(6, 2), _pad=True)),
    ("leftpad", (2, 16, 64, 2, 1, 32, 32), torch.bfloat16, dict(causal=True, _pad=True, _leftpad=True)),
    ("fp8_descale", (2, 32, 48, 4, 2, 64, 64), torch.bfloat16,
     dict(causal=True, _descale=True, upcast=False, reorder_ops=True, intermediate_dtype=torch.float8_e4m3fn)),
    ("fp8_descale_fp32", (2, 32, 48, 4, 2, 64, 64), torch.bfloat16, dict(causal=True, _descale=True)),
    # qv (scores += qv v^T, default scale 1/sqrt(d + dv)): appended after the round-1 cases so their seeds do not move
    ("qv_causal_gqa", (2, 24, 40, 4, 2, 64, 256), torch.bfloat16, dict(causal=True, _qv=True)),
    ("qv_lowprec", (1, 16, 33, 2, 1, 64, 256), torch.bfloat16, dict(_qv=True, upcast=False, reorder_ops=True)),
]


def main():
    out = {}
    for i, (name, (b, sq, sk, h, hk, d, dv), dtype, kw) in enumerate(CASES):
        g = torch.Generator().manual_seed(1000 + i)
        q = torch.randn(b, sq, h, d, generator=g).to(dtype)
        k = torch.randn(b, sk, hk, d, generator=g).to(dtype)
        v = torch.randn(b, sk, hk, dv, generator=g).to(dtype)
        kw = dict(kw)
        rec = dict(q=q, k=k, v=v, dtype=str(dtype))
        if kw.pop("_qv", False):
            kw["qv"] = torch.randn(b, sq, h, dv, generator=g).to(dtype)
            rec["qv"] = kw["qv"]
        if kw.pop("_descale", False):
            q = q.to(torch.float8_e4m3fn).to(dtype)
            k = k.to(torch.float8_e4m3fn).to(dtype)
            v = v.to(torch.float8_e4m3fn).to(dtype)
            rec.update(q=q, k=k, v=v)
            for nm in ("q_descale", "k_descale", "v_descale"):
                kw[nm] = torch.rand(b, hk, generator=g) * 2
        if kw.pop("_pad", False):
            qlen = torch.randint(1, sq + 1, (b,), generator=g)
            leftpad = None
            if kw.pop("_leftpad", False):
                leftpad = torch.randint(0, sk // 4, (b,), generator=g)
                kw["key_leftpad"] = leftpad
            klen = torch.randint(1, sk + 1 - (sk // 4 if leftpad is not None else 0), (b,), generator=g)
            qlen[0] = sq
            ar_q = torch.arange(sq).view(1, sq)
            ar_k = torch.arange(sk).view(1, sk)
            kw["query_padding_mask"] = ar_q < qlen.view(b, 1)
            lo = leftpad.view(b, 1) if leftpad is not None else 0
            kw["key_padding_mask"] = (ar_k >= lo) & (ar_k < lo + klen.view(b, 1))
            rec.update(q_lens=qlen, k_lens=klen)
        o, attn = attention_ref(q, k, v, **kw)
        rec["kwargs"] = {kk: vv for kk, vv in kw.items() if kk not in ("query_padding_mask", "key_padding_mask")}
        rec["out"] = o
        # row sums of the reference's probabilities: 1 for live rows, 0 for dead rows
        rec["attn_rowsum"] = attn.float().sum(-1)
        out[name] = rec

    # combine reference, hopper/test_flash_attn.py:1105-1114 evaluated verbatim in math
    g = torch.Generator().manual_seed(7)
    for name, (splits, b, s, h, d) in {"combine_a": (5, 3, 7, 4, 32), "combine_b": (17, 2, 33, 2, 64)}.items():
        op = torch.randn(splits, b, s, h, d, generator=g)
        lp = torch.randn(splits, b, s, h, generator=g)  # as the test builds it (:1139)
        lp[splits // 2:, : max(1, b // 3)] = float("-inf")
        lse = torch.logsumexp(lp, dim=0)
        sc = torch.exp(lp - lse)
        sc = torch.where(torch.isinf(sc) | torch.isnan(sc), torch.zeros_like(sc), sc)
        o = (sc.unsqueeze(-1) * op).sum(0)
        out[name] = dict(out_partial=op, lse_partial=lp, out=o, lse=lse)
    torch.save(out, os.path.join(HERE, "attention_ref.pt"))
    print("wrote", len(out), "cases,", os.path.getsize(os.path.join(HERE, "attention_ref.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
