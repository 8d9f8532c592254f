"""fa100_fwd_splitkv_noalloc (include/fa100.h): split-KV forward + combine with the fp32 partials in the library's static
device array -- no counterpart in the reference, added for callers that cannot
allocate device memory.  Called straight through the C ABI (ctypes) and checked against the existing entry points
(fa100_fwd with caller-supplied partials + fa100_combine, reached through fa100_ops.mha_fwd) on two calls in a row, with the
tolerance the other GPU tests use (2x the low-precision oracle's error + atol for O, 1e-3 relative for the LSE).

Named test_zz_* so that pytest collects it last: a failure here must not hide the
other tests.  Passes alone in a fresh process and after the rest of the suite (the scratch array keeps stale contents)."""
import ctypes as C

import pytest
import torch

import attention_oracle as A

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _params(N, q, k, v, out, lse, seqused_k, causal, num_splits, pack_gqa, num_sm):
    """Fa100FwdParams for dense (b, s, h, d) tensors, the way fa100_ops.mha_fwd fills it."""
    p = N.FwdParams()
    ptr = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731
    p.q, p.k, p.v, p.o, p.lse = ptr(q), ptr(k), ptr(v), ptr(out), ptr(lse)
    p.q_batch_stride, p.q_row_stride, p.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
    p.k_batch_stride, p.k_row_stride, p.k_head_stride = k.stride(0), k.stride(1), k.stride(2)
    p.v_batch_stride, p.v_row_stride, p.v_head_stride = v.stride(0), v.stride(1), v.stride(2)
    p.o_batch_stride, p.o_row_stride, p.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    b, sq, h, d = q.shape
    p.b, p.b_k, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d, p.dv = b, k.shape[0], sq, k.shape[1], h, k.shape[2], d, v.shape[3]
    p.dtype = N.BF16 if q.dtype == torch.bfloat16 else N.FP16
    p.seqused_k = ptr(seqused_k)
    p.softmax_scale, p.softcap = float(d ** -0.5), 0.0
    p.is_causal, p.window_left, p.window_right, p.attention_chunk = int(causal), -1, (0 if causal else -1), 0
    p.num_splits, p.pack_gqa, p.num_sm = num_splits, pack_gqa, num_sm
    return p


def _run_noalloc(q, k, v, lens, causal, num_splits):
    import fa100_native as N
    import fa100_ops as ops

    b, sq, h, d = q.shape
    pl = ops.plan(b, sq, k.shape[1], h, k.shape[2], d, v.shape[3], q.dtype, causal, False, num_splits, None, 0, q.device)
    out = torch.empty((b, sq, h, v.shape[3]), dtype=q.dtype, device=q.device)
    lse = torch.empty((b, h, sq), dtype=torch.float32, device=q.device)
    p = _params(N, q, k, v, out, lse, lens, causal, pl.num_splits, pl.pack_gqa, pl.num_sm)
    before = torch.cuda.memory_allocated()
    N.check(N.lib().fa100_fwd_splitkv_noalloc(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    assert torch.cuda.memory_allocated() == before  # nothing was allocated through torch on the way
    return out, lse


def _reference_path(q, k, v, lens, causal, num_splits):
    """The existing entry points: fa100_fwd with partials the caller allocates, then fa100_combine."""
    import fa100_ops as ops

    out, lse, *_ = ops.mha_fwd(q, k, v, seqused_k=lens, is_causal=causal, num_splits=num_splits)
    return out, lse


def _compare(case, q, k, v, lens, causal, num_splits):
    import fa100_native as N

    got_o, got_l = _run_noalloc(q, k, v, lens, causal, num_splits)
    want_o, want_l = _reference_path(q, k, v, lens, causal, num_splits)
    torch.cuda.synchronize()
    assert N.hang_info()[0] == 0
    kl = None if lens is None else lens.cpu()
    ref, lse_ref = A.attention_oracle(q.cpu(), k.cpu(), v.cpu(), k_lens=kl, causal=causal)
    low, _ = A.attention_oracle(q.cpu(), k.cpu(), v.cpu(), k_lens=kl, causal=causal, upcast=False, reorder_ops=True)
    tol = A.fwd_tolerance(ref, low, 2.0)
    assert torch.isfinite(got_o.float()).all(), case
    assert (got_o.float() - want_o.float()).abs().max().item() <= tol, case          # new entry point vs the existing one
    assert (got_o.float().cpu() - ref.float()).abs().max().item() <= tol, case       # and vs the oracle, same rule
    assert torch.equal(torch.isneginf(got_l), torch.isneginf(want_l)), case
    fin = torch.isfinite(want_l)
    assert ((got_l[fin] - want_l[fin]).abs() / want_l[fin].abs().clamp_min(1.0)).max().item() <= 1e-3, case
    fin = torch.isfinite(lse_ref)
    assert ((got_l.cpu()[fin] - lse_ref[fin]).abs() / lse_ref[fin].abs().clamp_min(1.0)).max().item() <= 1e-3, case


def test_noalloc_splitkv_matches_existing_entry_point_twice():
    g = torch.Generator(device=DEV).manual_seed(91)
    mk = lambda *s: torch.randn(*s, generator=g, device=DEV, dtype=torch.bfloat16)  # noqa: E731
    # call 1: decode, 8 sequences of up to 2048 keys, 4 partials (dual-stream decode kernel)
    q, k, v = mk(8, 1, 16, 128), mk(8, 2048, 4, 128), mk(8, 2048, 4, 128)
    lens = torch.tensor([2048, 1500, 129, 1, 777, 2000, 64, 1024], dtype=torch.int32, device=DEV)
    _compare("decode, 4 partials", q, k, v, lens, False, 4)
    # call 2, straight after, other sizes and fewer partials: stale contents of the shared scratch must not leak in
    q2, k2, v2 = mk(2, 64, 8, 128), mk(2, 1024, 2, 128), mk(2, 1024, 2, 128)
    lens2 = torch.tensor([1024, 300], dtype=torch.int32, device=DEV)
    _compare("sq 64 causal, 3 partials", q2, k2, v2, lens2, True, 3)
    # call 3 = call 1 again: same result as before after the array was reused
    _compare("decode again", q, k, v, lens, False, 4)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import fa100_native as N

    assert N.lib().fa100_noalloc_scratch_bytes() == 16 * 1024 * 1024
    q = torch.empty((4, 1024, 32, 128), dtype=torch.bfloat16, device=DEV)
    k = torch.empty((4, 2048, 8, 128), dtype=torch.bfloat16, device=DEV)
    out = torch.empty_like(q)
    lse = torch.empty((4, 32, 1024), dtype=torch.float32, device=DEV)
    p = _params(N, q, k, k, out, lse, None, False, 4, 0, 148)  # 4 x 4 x 32 x 1024 x 129 x 4 B = 270 MB of partials
    rc = N.lib().fa100_fwd_splitkv_noalloc(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == N.ERR_INVALID
    assert b"scratch too small" in N.lib().fa100_last_error()
    p.num_splits = 1
    assert N.lib().fa100_fwd_splitkv_noalloc(C.byref(p), None) == N.ERR_INVALID  # a single pass has no scratch: use fa100_fwd
    torch.cuda.synchronize()
