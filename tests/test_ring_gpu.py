"""Ring (context-parallel) attention on the GPU: the CUDA backend (fa100_fwd partial mode + fa100_combine) driven by
the ring schedule.  Single device: every rank of a simulated 2/4/8-way ring runs in this process.  Two devices (skipped
when only one is visible): a real NCCL ring, launched as two processes."""
import os
import sys

import pytest
import torch

import attention_oracle as A

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _inputs(b, s, h, hk, d, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mk = lambda *shape: torch.randn(*shape, generator=g, device="cuda", dtype=torch.bfloat16)  # noqa: E731
    return mk(b, s, h, d), mk(b, s, hk, d), mk(b, s, hk, d)


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("s,h,hk", [(2048, 4, 4), (4096, 8, 2)])
def test_simulated_ring_matches_single_pass(world, s, h, hk):
    import ring_attention as R
    from flash_attn_interface import flash_attn_func

    q, k, v = _inputs(2, s, h, hk, 128, 5)
    out, lse = R.simulate_ring(q, k, v, world)
    ref_o, ref_lse = flash_attn_func(q, k, v, causal=True, return_attn_probs=True)
    # oracle on a slice (fp32), reference tolerance rule; and agreement with the single-pass kernel
    oref, lref = A.attention_oracle(q[:1, :, :h // hk].cpu(), k[:1, :, :1].cpu(), v[:1, :, :1].cpu(), causal=True)
    low, _ = A.attention_oracle(q[:1, :, :h // hk].cpu(), k[:1, :, :1].cpu(), v[:1, :, :1].cpu(), causal=True, upcast=False, reorder_ops=True)
    err = (out[:1, :, :h // hk].float().cpu() - oref.float()).abs().max().item()
    assert err <= A.fwd_tolerance(oref, low, 2.0), err
    assert (out.float() - ref_o.float()).abs().max().item() <= 2e-2
    assert torch.allclose(lse, ref_lse, atol=1e-4, rtol=1e-4)
    assert torch.allclose(lse[:1, :h // hk].cpu(), lref, atol=1e-3, rtol=1e-3)


def _nccl_worker(rank, world, port, q_out, transport="nccl"):
    sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200"), os.path.join(ROOT, "oracle")]
    import torch.distributed as dist
    import ring_attention as R
    from flash_attn_interface import flash_attn_func

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        g = torch.Generator(device="cuda").manual_seed(77)  # same full problem on every rank
        mk = lambda *shape: torch.randn(*shape, generator=g, device="cuda", dtype=torch.bfloat16)  # noqa: E731
        q, k, v = mk(1, 8192, 8, 128), mk(1, 8192, 2, 128), mk(1, 8192, 2, 128)  # GQA shards: K / V messages smaller than Q
        ref_o, ref_lse = flash_attn_func(q, k, v, causal=True, return_attn_probs=True)
        err = lerr = 0.0
        for _ in range(3):  # repeated calls reuse the transport's buffers (flow control across calls)
            out, lse = R.ring_flash_attn_func(R.zigzag_shard(q, rank, world), R.zigzag_shard(k, rank, world),
                                              R.zigzag_shard(v, rank, world), causal=True, return_lse=True, transport=transport,
                                              sm_margin=8 if transport == "nccl" else 0)
            err = max(err, (out.float() - R.zigzag_shard(ref_o, rank, world).float()).abs().max().item())
            lerr = max(lerr, (lse - R.zigzag_shard(ref_lse, rank, world, dim=2)).abs().max().item())
        torch.cuda.synchronize()
        q_out.put((rank, err, lerr))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (run with gpurun --gpus 2)")
@pytest.mark.parametrize("transport", ["nccl", "ce_push"])
def test_nccl_ring_two_gpus(transport):
    """The real ring over NVLink on 2 (or 4 / 8 when visible) GPUs, one process per GPU: NCCL send/recv ring, and the
    copy-engine push transport (peer memory + stream memory operations, no communication kernels)."""
    import torch.multiprocessing as mp

    world = torch.cuda.device_count()
    world = 8 if world >= 8 else (4 if world >= 4 else 2)
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, 29871 + (transport == "ce_push"), q_out, transport)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, lerr in res:
        assert err <= 2e-2 and lerr <= 1e-4, (rank, err, lerr)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_fused_merge_equals_partial_plus_combine(dtype):
    """Fa100FwdParams.merge_partial: the epilogue's merge (1 = in place, 2 = into the 16-bit output) against the same two
    partials merged by fa100_combine (hopper/flash_fwd_combine_kernel.h:345-458); rows with no keys on either side."""
    import fa100_ops as ops

    b, sq, h, hk, d = 2, 300, 4, 2, 128
    g = torch.Generator(device="cuda").manual_seed(11)
    mk = lambda *shape: torch.randn(*shape, generator=g, device="cuda", dtype=torch.float32).to(dtype)  # noqa: E731
    q, k1, v1, k2, v2 = mk(b, sq, h, d), mk(b, 500, hk, d), mk(b, 500, hk, d), mk(b, 260, hk, d), mk(b, 260, hk, d)
    o = torch.empty((2, b, h, sq, d), dtype=torch.float32, device="cuda")
    l = torch.empty((2, b, h, sq), dtype=torch.float32, device="cuda")
    ops.mha_fwd(q, k1, v1, is_causal=True, _partial=(o[:1], l[:1]))   # bottom-right causal: the first rows see few keys
    ops.mha_fwd(q, k2, v2, is_causal=True, _partial=(o[1:], l[1:]))   # sk < sq: the first 40 rows see no key at all
    want_o = torch.empty((b, sq, h, d), dtype=dtype, device="cuda")
    want_l = torch.empty((b, sq, h), dtype=torch.float32, device="cuda")
    ops._combine_into(o.permute(0, 1, 3, 2, 4), l.permute(0, 1, 3, 2), want_o, want_l)
    acc_o, acc_l = o[:1].clone(), l[:1].clone()
    ops.mha_fwd(q, k2, v2, is_causal=True, _partial=(acc_o, acc_l), _merge=1)
    got_o = torch.empty((b, sq, h, d), dtype=dtype, device="cuda")
    acc2_o, acc2_l = o[:1].clone(), l[:1].clone()
    _, got_l, *_ = ops.mha_fwd(q, k2, v2, out=got_o, is_causal=True, _partial=(acc2_o, acc2_l), _merge=2)
    torch.cuda.synchronize()
    assert torch.isneginf(l[1, :, :, :40]).all() and torch.isfinite(want_l).all()
    fp32_want = torch.empty((b, sq, h, d), dtype=torch.float32, device="cuda")
    ops._combine_into(o.permute(0, 1, 3, 2, 4), l.permute(0, 1, 3, 2), fp32_want, torch.empty_like(want_l))
    assert torch.allclose(acc_o[0].permute(0, 2, 1, 3), fp32_want, atol=2e-5, rtol=1e-5)
    assert torch.allclose(acc_l[0].permute(0, 2, 1), want_l, atol=1e-5, rtol=1e-5)
    assert (got_o.float() - want_o.float()).abs().max().item() <= (8e-3 if dtype == torch.bfloat16 else 1e-3)
    assert torch.allclose(got_l.permute(0, 2, 1), want_l, atol=1e-5, rtol=1e-5)
