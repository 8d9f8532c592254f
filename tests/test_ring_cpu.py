"""Ring (context-parallel) attention host logic on CPU: world_size 2 and 4 over gloo, arithmetic supplied by the
oracle (tests may use it; the product backend is CUDA only).  Checks that the zig-zag schedule, the per-hop masks,
the merge order and the K/V exchange reproduce single-device causal attention exactly (fp32 merge, 1e-5)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import attention_oracle as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleBackend:
    """Same interface as ring_attention.Fa100Backend, computed with the CPU oracle (test double)."""

    def alloc(self, b, h, rows, dv, device):
        return torch.zeros(2, b, h, rows, dv), torch.zeros(2, b, h, rows)

    def partial(self, q, k, v, causal, o_slot, lse_slot):
        o, lse = A.attention_oracle(q.float(), k.float(), v.float(), causal=causal)
        o_slot.copy_(o.permute(0, 2, 1, 3))
        lse_slot.copy_(lse)

    def merge(self, o, lse, out=None, out_lse=None):
        mo, ml = A.combine_oracle(o.permute(0, 1, 3, 2, 4), lse.permute(0, 1, 3, 2))  # (b, rows, h, dv), (b, rows, h)
        if out is None:
            o[0].copy_(mo.permute(0, 2, 1, 3))
            lse[0].copy_(ml.permute(0, 2, 1))
        else:
            out.copy_(mo)
            out_lse.copy_(ml.permute(0, 2, 1))

    def finish_single(self, o, lse, out, out_lse):
        out.copy_(o[0].permute(0, 2, 1, 3))
        out_lse.copy_(lse[0])


def _worker(rank, world, port, seq, h, hk, d, q_out):
    sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200"), os.path.join(ROOT, "oracle")]
    import ring_attention as R

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(123)
        q = torch.randn(2, seq, h, d, generator=g)
        k = torch.randn(2, seq, hk, d, generator=g)
        v = torch.randn(2, seq, hk, d, generator=g)
        out, lse = R.ring_flash_attn_func(R.zigzag_shard(q, rank, world), R.zigzag_shard(k, rank, world), R.zigzag_shard(v, rank, world),
                                          causal=True, backend=OracleBackend(), return_lse=True)
        ref, lse_ref = A.attention_oracle(q, k, v, causal=True)
        err = (out - R.zigzag_shard(ref, rank, world)).abs().max().item()
        lerr = (lse - R.zigzag_shard(lse_ref, rank, world, dim=2)).abs().max().item()
        q_out.put((rank, err, lerr))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,seq", [(2, 64), (4, 128), (3, 60)])
def test_ring_matches_single_device(world, seq):
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = 29600 + world * 7 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, world, port, seq, 4, 2, 32, q_out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, lerr in res:
        assert err < 1e-5 and lerr < 1e-5, (rank, err, lerr)


def test_zigzag_roundtrip_and_plan():
    sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
    import ring_attention as R

    x = torch.arange(2 * 48 * 3).view(2, 48, 3)
    for world in (1, 2, 3, 4):
        shards = [R.zigzag_shard(x, r, world) for r in range(world)]
        assert torch.equal(R.zigzag_unshard(shards), x)
        # every (row chunk, key chunk) pair of the causal triangle is covered exactly once across hops and ranks
        seen = set()
        for r in range(world):
            for t, (s, work) in enumerate(R.hop_plan(r, world)):
                assert s == (r - t) % world
                for g, keys, causal in work:
                    rc = r if g == "lo" else 2 * world - 1 - r
                    for kc in ([s] if keys == "lo" else [s, 2 * world - 1 - s]):
                        assert (rc, kc) not in seen and kc <= rc and (causal == (kc == rc) or (causal and kc < rc and keys == "all"))
                        seen.add((rc, kc))
        assert seen == {(a, b) for a in range(2 * world) for b in range(a + 1)}
