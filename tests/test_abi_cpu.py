"""CPU: the C-ABI library builds, loads and exports every symbol include/fa100.h declares; ctypes mirrors of the
structs have the C sizes; the sizes-only planner and argument validation work without a GPU; the op schemas match
the reference's frozen ones (hopper/test_flash_attn.py:1163-1201)."""
import ctypes as C
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def native():
    import __graft_entry__ as G

    G.build()
    import fa100_native as N

    return N


def test_exports_match_header(native):
    hdr = open(os.path.join(ROOT, "include", "fa100.h")).read()
    declared = set(re.findall(r"^(?:int|int64_t|const char\*)\s+(fa100_\w+)\s*\(", hdr, flags=re.M))
    assert declared == set(native.EXPORTS)
    lib = native.lib()
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.fa100_abi_version() == 2  # 2: Fa100FwdParams grew the qv fields


def test_struct_sizes_match_c(native, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "fa100.h"\nint main(){printf("%zu %zu %zu %zu %zu\\n", sizeof(Fa100FwdParams),'
                   'sizeof(Fa100CombineParams), sizeof(Fa100Shape), sizeof(Fa100Plan), sizeof(Fa100BwdParams));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert sizes == [C.sizeof(native.FwdParams), C.sizeof(native.CombineParams), C.sizeof(native.Shape), C.sizeof(native.Plan), C.sizeof(native.BwdParams)]


def _plan(native, **kw):
    base = dict(b=1, seqlen_q=1, seqlen_k=1, h=1, h_k=1, d=128, dv=128, dtype=0, num_splits=0, pack_gqa=-1, num_sm=148)
    base.update(kw)
    s, p = native.Shape(**base), native.Plan()
    rc = native.lib().fa100_plan(C.byref(s), C.byref(p))
    return rc, p


def test_plan_policy(native):
    rc, p = _plan(native, b=8, seqlen_q=8192, seqlen_k=8192, h=32, h_k=8, is_causal=1)
    assert rc == 0 and p.num_splits == 1 and p.pack_gqa == 0 and p.num_work == 8 * 32 * 32
    assert p.dual_stream == 0                                                      # two live Q tiles per item: single stream
    rc, p = _plan(native, b=64, seqlen_q=1, seqlen_k=32768, h=32, h_k=8)          # C4: decode packs the 4 group heads and
    assert rc == 0 and p.pack_gqa == 1 and p.dual_stream == 1                      # runs dual-stream: two partials per CTA split
    assert p.num_splits == 2 and p.num_work == 64 * 8 * (p.num_splits // 2)
    rc, p = _plan(native, b=64, seqlen_q=1, seqlen_k=32768, h=32, h_k=8, num_splits=3)   # a forced odd split count keeps
    assert rc == 0 and p.dual_stream == 0 and p.num_splits == 3 and p.num_work == 64 * 8 * 3  # the single-stream kernel
    rc, p = _plan(native, b=1, seqlen_q=1, seqlen_k=32768, h=8, h_k=8)             # tiny batch: split to fill the SMs
    assert rc == 0 and p.num_splits > 8 and p.num_splits % 2 == 0 and p.dual_stream == 1 and p.num_work <= 148
    rc, p = _plan(native, b=2, seqlen_q=64, seqlen_k=100, h=4, h_k=4)              # one key block: never split
    assert rc == 0 and p.num_splits == 1
    rc, p = _plan(native, b=2, seqlen_q=64, seqlen_k=4096, h=4, h_k=4, num_splits=7, pack_gqa=0)  # forced values are kept
    assert rc == 0 and p.num_splits == 7 and p.pack_gqa == 0
    rc, p = _plan(native, b=4, seqlen_q=8192, seqlen_k=8192, h=16, h_k=4, d=256, dv=256, is_causal=1)   # head dim > 128:
    assert rc == 0 and p.block_m == 128 and p.num_work == 4 * 16 * 64 and p.dual_stream == 0            # one Q tile per item
    rc, p = _plan(native, b=64, seqlen_q=1, seqlen_k=32768, h=32, h_k=8, d=192, dv=128)                  # and never dual-stream
    assert rc == 0 and p.dual_stream == 0 and p.block_m == 128
    rc, _ = _plan(native, h=6, h_k=4)
    assert rc == native.ERR_INVALID and b"multiple" in native.lib().fa100_last_error()


def test_argument_validation_without_gpu(native):
    p = native.FwdParams()
    assert native.lib().fa100_fwd(C.byref(p), None) != 0          # null tensors / zero sizes are rejected before any launch
    c = native.CombineParams()
    assert native.lib().fa100_combine(C.byref(c), None) == native.ERR_INVALID
    assert native.lib().fa100_fwd(None, None) == native.ERR_INVALID
    with pytest.raises(RuntimeError):
        native.check(native.ERR_INVALID)
    with pytest.raises(NotImplementedError):
        native.check(native.ERR_UNSUPPORTED)


def test_op_schemas_match_reference(native):
    """The WHOLE schema string of every op equals the reference's (hopper/flash_api.cpp:1672-1761, frozen by
    hopper/test_flash_attn.py:1163-1201); tests/golden/op_schemas.json holds torch's canonical spelling of the reference's
    m.def literals (tests/golden/make_golden_schemas.py)."""
    import json

    import flash_attn_interface  # noqa: F401

    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "op_schemas.json")))
    assert set(gold) == {"fwd", "bwd", "fwd_combine", "get_scheduler_metadata"}
    for name, want in gold.items():
        assert str(getattr(torch.ops.flash_attn_3, name).default._schema) == want, name


def test_public_signatures_match_reference(native):
    """argument names and defaults of the public functions (hopper/flash_attn_interface.py:507-523,588-610,640-671,803-820)."""
    import inspect

    import flash_attn_interface as F

    sig = lambda f: [(n, p.default) for n, p in inspect.signature(f).parameters.items()]  # noqa: E731
    E = inspect.Parameter.empty
    assert sig(F.flash_attn_func) == [("q", E), ("k", E), ("v", E), ("softmax_scale", None), ("causal", False), ("qv", None),
                                      ("q_descale", None), ("k_descale", None), ("v_descale", None), ("window_size", (-1, -1)),
                                      ("attention_chunk", 0), ("softcap", 0.0), ("num_splits", 1), ("pack_gqa", None),
                                      ("deterministic", False), ("sm_margin", 0), ("return_attn_probs", False)]
    names = [n for n, _ in sig(F.flash_attn_varlen_func)]
    assert names[:9] == ["q", "k", "v", "cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k", "seqused_q", "seqused_k"]
    kv = dict(sig(F.flash_attn_with_kvcache))
    assert list(kv)[:9] == ["q", "k_cache", "v_cache", "k", "v", "qv", "rotary_cos", "rotary_sin", "cache_seqlens"]
    assert kv["num_splits"] == 0 and kv["rotary_interleaved"] is True and kv["return_softmax_lse"] is False
    assert [n for n, _ in sig(F.flash_attn_combine)] == ["out_partial", "lse_partial", "out", "out_dtype"]
    assert [n for n, _ in sig(F.get_scheduler_metadata)][:7] == ["batch_size", "max_seqlen_q", "max_seqlen_k", "num_heads_q",
                                                                 "num_heads_kv", "headdim", "cache_seqlens"]


def test_no_cpu_fallback(native):
    import flash_attn_interface as F

    q = torch.randn(1, 8, 2, 64, dtype=torch.bfloat16)
    with pytest.raises((RuntimeError, NotImplementedError)):
        F.flash_attn_func(q, q, q)


def test_meta_kernels_trace_shapes():
    """FakeTensor / Meta dispatch: flash_attn_3::fwd and ::fwd_combine infer output shapes without a GPU
    (what torch.compile needs to trace through the ops)."""
    import torch

    import flash_attn_interface  # noqa: F401  (registers the ops)

    q = torch.empty(2, 300, 8, 128, dtype=torch.bfloat16, device="meta")
    k = torch.empty(2, 500, 2, 128, dtype=torch.bfloat16, device="meta")
    v = torch.empty(2, 500, 2, 64, dtype=torch.bfloat16, device="meta")
    out, lse, oacc, lacc = torch.ops.flash_attn_3.fwd(q, k, v, is_causal=True)
    assert out.shape == (2, 300, 8, 64) and out.dtype == torch.bfloat16 and lse.shape == (2, 8, 300) and lse.dtype == torch.float32
    assert oacc.numel() == 0 and lacc.numel() == 0
    out, lse, oacc, lacc = torch.ops.flash_attn_3.fwd(q, k, v, num_splits=3)
    assert oacc.shape == (3, 2, 8, 300, 64) and lacc.shape == (3, 2, 8, 300)
    q8 = torch.empty(700, 4, 128, dtype=torch.float8_e4m3fn, device="meta")
    cu = torch.empty(4, dtype=torch.int32, device="meta")
    out, lse, *_ = torch.ops.flash_attn_3.fwd(q8, q8, q8, cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=300, max_seqlen_k=300)
    assert out.shape == (700, 4, 128) and out.dtype == torch.bfloat16 and lse.shape == (4, 700)
    op = torch.empty(5, 2, 100, 4, 64, device="meta")
    lp = torch.empty(5, 2, 100, 4, device="meta")
    o, l = torch.ops.flash_attn_3.fwd_combine(op, lp, None, torch.bfloat16)
    assert o.shape == (2, 100, 4, 64) and o.dtype == torch.bfloat16 and l.shape == (2, 100, 4)


def test_fa2_compat_signatures_match_reference():
    """Names, order and defaults of the FA2-style entry points (compat/flash_attn/flash_attn_interface.py) equal the
    reference's flash_attn/flash_attn_interface.py:1003-1590 (tests/golden/fa2_signatures.json, extracted with ast)."""
    import ast
    import json

    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fa2_signatures.json")))
    src = open(os.path.join(ROOT, "flash-attention-annotated_b200", "compat", "flash_attn", "flash_attn_interface.py")).read()
    mine = {}
    for n in ast.parse(src).body:
        if isinstance(n, ast.FunctionDef) and n.name.startswith("flash_attn"):
            a = n.args
            names = [x.arg for x in a.args]
            d = [ast.unparse(x) for x in a.defaults]
            mine[n.name] = [[x, y] for x, y in zip(names, [None] * (len(names) - len(d)) + d)]
    assert set(mine) == set(gold)
    for name in gold:
        assert mine[name] == gold[name], name
