"""Context-parallel (ring) causal attention over the fa100 forward -- BASELINE config C5, SURVEY.md 8(e).

The reference has no multi-GPU attention; its contract for merging partial results is the combine kernel
(hopper/flash_fwd_combine_kernel.h:345-458, oracle ``attention_combine_ref`` hopper/test_flash_attn.py:1105-1114):
``LSE = logaddexp_s LSE_s``, ``O = sum_s exp(LSE_s - LSE) O_s``.  This module shards ONE long sequence over the ranks of
a process group and computes exactly ``flash_attn_func(q, k, v, causal=True)`` of the unsharded problem.

Layout (zig-zag): the sequence is cut into 2N chunks of c rows; rank r owns chunks r ("lo") and 2N-1-r ("hi") of Q, K
and V, so every rank has the same causal work in every hop.  Hop t uses the K/V that originated at rank s = (r-t) mod N:

  t = 0 (s = r)   lo rows x lo keys causal;  hi rows x (lo, hi) keys causal (bottom-right aligned)
  s < r           lo rows and hi rows x lo keys of s, unmasked   (hi keys of s lie in the future of both)
  s > r           hi rows x (lo, hi) keys of s, unmasked         (lo rows see nothing of s)

The first launch of a row group writes an fp32 partial (O, LSE); every later hop's launch merges its own result into it in
the attention kernel's epilogue (``Fa100FwdParams.merge_partial``), the last one writing the 16-bit output -- no separate
combine launch and no second partial buffer on the CUDA path (the CPU test backend keeps the two-way ``fa100_combine``
form of the same arithmetic).  K/V travel around the ring as one
packed buffer with NCCL send/recv on a side stream, double-buffered so hop t+1's transfer overlaps hop t's attention;
the final O stays sequence-sharded (no gather).  Message per hop: 2 * 2c * h_k * d * 2 B (268 MB at C5 on 8 GPUs).

The arithmetic lives behind a small backend object so that the ring logic (schedule, masks, merge order, exchange)
is testable on CPU with gloo; the product backend is CUDA only and there is no fallback.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


# --------------------------------------------------------------------------- zig-zag sharding helpers
def zigzag_shard(x: torch.Tensor, rank: int, world: int, dim: int = 1) -> torch.Tensor:
    """Rows of rank `rank`: chunks rank and 2*world-1-rank of `x` along `dim`, concatenated (lo then hi)."""
    n = x.shape[dim]
    assert n % (2 * world) == 0, "sequence length must be a multiple of 2 * world_size"
    chunks = x.chunk(2 * world, dim=dim)
    return torch.cat((chunks[rank], chunks[2 * world - 1 - rank]), dim=dim).contiguous()


def zigzag_unshard(shards: List[torch.Tensor], dim: int = 1) -> torch.Tensor:
    """Inverse of zigzag_shard given the shards of ranks 0..world-1."""
    world = len(shards)
    chunks: List[Optional[torch.Tensor]] = [None] * (2 * world)
    for r, s in enumerate(shards):
        lo, hi = s.chunk(2, dim=dim)
        chunks[r], chunks[2 * world - 1 - r] = lo, hi
    return torch.cat(chunks, dim=dim)


def hop_plan(rank: int, world: int) -> List[Tuple[int, List[Tuple[str, str, bool]]]]:
    """[(source rank, [(row group, key range, causal), ...]), ...] for hops t = 0..world-1.
    row group: "lo" | "hi"; key range: "lo" (first c keys of the source's buffer) | "all" (its 2c keys)."""
    plan = []
    for t in range(world):
        s = (rank - t) % world
        if t == 0:
            plan.append((s, [("lo", "lo", True), ("hi", "all", True)]))
        elif s < rank:
            plan.append((s, [("lo", "lo", False), ("hi", "lo", False)]))
        else:
            plan.append((s, [("hi", "all", False)]))
    return plan


# --------------------------------------------------------------------------- CUDA backend (product path)
class Fa100Backend:
    """Partial attention = fa100_fwd in partial mode; merge = fa100_combine.  CUDA only."""

    def __init__(self, softmax_scale: Optional[float] = None, sm_margin: int = 0):
        import fa100_ops as ops  # raises if the library is missing; no fallback

        self.ops = ops
        self.scale = softmax_scale
        self.sm_margin = sm_margin  # SMs left to the NCCL send/recv kernels while a transfer overlaps the hop

    def alloc(self, b, h, rows, dv, device):
        o = torch.empty((1, b, h, rows, dv), dtype=torch.float32, device=device)  # the running partial (later hops merge into it in the kernel)
        lse = torch.empty((1, b, h, rows), dtype=torch.float32, device=device)
        return o, lse

    fused_merge = True  # the attention epilogue merges into the running partial itself: no combine launch

    def partial(self, q, k, v, causal, o_slot, lse_slot):
        self.ops.mha_fwd(q, k, v, softmax_scale=self.scale, is_causal=causal, sm_margin=self.sm_margin, _partial=(o_slot.unsqueeze(0), lse_slot.unsqueeze(0)))

    def partial_merge(self, q, k, v, causal, o_slot, lse_slot, out=None, out_lse=None):
        """This hop's attention merged into the running fp32 partial in `o_slot` / `lse_slot` by the kernel's epilogue; the
        last hop (`out` given) writes the 16-bit output (b, rows, h, dv) and the LSE (b, h, rows) instead."""
        if out is None:
            self.ops.mha_fwd(q, k, v, softmax_scale=self.scale, is_causal=causal, sm_margin=self.sm_margin, _partial=(o_slot.unsqueeze(0), lse_slot.unsqueeze(0)), _merge=1)
        else:
            _, lse, *_ = self.ops.mha_fwd(q, k, v, out=out, softmax_scale=self.scale, is_causal=causal, sm_margin=self.sm_margin,
                                          _partial=(o_slot.unsqueeze(0), lse_slot.unsqueeze(0)), _merge=2)
            out_lse.copy_(lse)

    def merge(self, o, lse, out=None, out_lse=None):
        """slot 0 <- merge(slot 0, slot 1) in fp32, or into the 16-bit `out` (b, rows, h, dv) / `out_lse` (b, h, rows)."""
        op = o.permute(0, 1, 3, 2, 4)   # (2, b, rows, h, dv)
        lp = lse.permute(0, 1, 3, 2)    # (2, b, rows, h)
        if out is None:
            self.ops._combine_into(op, lp, op[0], lp[0])
        else:
            self.ops._combine_into(op, lp, out, out_lse.permute(0, 2, 1))

    def finish_single(self, o, lse, out, out_lse):
        out.copy_(o[0].permute(0, 2, 1, 3))
        out_lse.copy_(lse[0])


# --------------------------------------------------------------------------- K/V exchange
class _RingExchange:
    """NCCL send/recv ring (north_star's transport).  K and V are two messages of one grouped send/recv, so d != dv and
    GQA shards need no packing; hop 0 uses the caller's tensors as they are (no staging copy), later hops alternate
    between two receive buffers so that hop t+1's transfer overlaps hop t's attention."""

    name = "nccl"

    def __init__(self, k, v, rank, world, group, events=None):
        self.cur = (k, v)
        self.bufs = [(torch.empty_like(k), torch.empty_like(v)) for _ in range(2)] if world > 1 else []
        self.world, self.group, self.dev = world, group, k.device
        self.comm_stream = torch.cuda.Stream(device=k.device) if (k.is_cuda and world > 1) else None
        self.nxt, self.prv = (rank + 1) % world, (rank - 1) % world
        if group is not None and world > 1:
            self.nxt, self.prv = dist.get_global_rank(group, self.nxt), dist.get_global_rank(group, self.prv)
        self.reqs, self.events = [], events

    def begin(self, t):
        """Start moving the buffer of hop t on to the next rank (and receiving hop t+1's); return hop t's (K, V)."""
        cur = self.cur
        if t + 1 < self.world:
            recv = self.bufs[(t + 1) % 2]
            ops = [dist.P2POp(dist.isend, cur[0], self.nxt, self.group), dist.P2POp(dist.isend, cur[1], self.nxt, self.group),
                   dist.P2POp(dist.irecv, recv[0], self.prv, self.group), dist.P2POp(dist.irecv, recv[1], self.prv, self.group)]
            if self.comm_stream is not None:
                self.comm_stream.wait_stream(torch.cuda.current_stream(self.dev))  # recv's last readers (hop t-1) are done
                with torch.cuda.stream(self.comm_stream):
                    e0 = _mark(self.events, self.comm_stream)
                    self.reqs = dist.batch_isend_irecv(ops)
                    for r in self.reqs:
                        r.wait()  # stream-level: the side stream waits for NCCL's own stream, the host does not block
                    e1 = _mark(self.events, self.comm_stream)
                if self.events is not None:
                    self.events.append(("comm", t, e0, e1))
            else:
                self.reqs = dist.batch_isend_irecv(ops)
                for r in self.reqs:
                    r.wait()
            self.nxt_cur = recv
        return cur

    def end(self, t):
        if self.comm_stream is not None and self.reqs:
            torch.cuda.current_stream(self.dev).wait_stream(self.comm_stream)
        if self.reqs:
            self.cur = self.nxt_cur
        self.reqs = []

    def finish(self):
        pass


_ring_groups = {}


def nccl_ring_group(sm_margin: int, ranks=None):
    """A dedicated NCCL communicator for the K/V ring whose kernels fit into the `sm_margin` SMs the attention kernel leaves
    free: `max_ctas = sm_margin` (ncclConfig_t.maxCTAs) and a high-priority stream.  The persistent attention kernel holds its
    SMs for the whole hop, so any NCCL channel CTA that does not fit beside it only runs once the hop is over -- measured on
    4 GPUs with the default communicator: 537 MB hops took 7.5 ms (the length of the hop) instead of ~2 ms.
    Collective: every rank of `ranks` (default: all) must call it with the same arguments.  Cached per (sm_margin, ranks)."""
    key = (int(sm_margin), None if ranks is None else tuple(ranks))
    if key not in _ring_groups:
        opts = dist.ProcessGroupNCCL.Options()
        opts.is_high_priority_stream = True
        opts.config.max_ctas = max(1, int(sm_margin))
        opts.config.min_ctas = 1
        _ring_groups[key] = dist.new_group(ranks=ranks, backend="nccl", pg_options=opts)
    return _ring_groups[key]


def _mark(events, stream):
    if events is None:
        return None
    e = torch.cuda.Event(enable_timing=True)
    e.record(stream)
    return e


class _PushExchange:
    """Copy-engine transport over NVSwitch peer memory: no SM is taken from the attention kernel.

    NVSwitch gives every GPU full bandwidth to every peer, so nothing has to be forwarded around a ring: each rank PUSHES
    its K/V shard straight into a receive slot in every peer's memory (symmetric-memory buffer, peer-mapped), in the order
    the peers need it (peer r+t uses it at its hop t).  The copies are cudaMemcpyAsync on a side stream -> DMA engines; the
    hand-shakes are stream memory operations (cuStreamWriteValue32 on the peer's signal pad after the copy,
    cuStreamWaitValue32 on the local pad in front of the hop), so neither side launches a kernel for communication and the
    persistent attention kernel keeps all 148 SMs.  The NCCL send/recv kernels of _RingExchange need SMs of their own while
    the attention kernel's 384 threads x 168 registers fill every SM it runs on.

    Flow control across calls: the receive slots are double-buffered by call parity, so before overwriting its slot on peer
    p a rank only waits until p has signalled "consumed" for the call BEFORE the previous one (written by p's compute stream
    after its last hop): consecutive calls of different ranks may drift apart by a whole call without stalling each other.
    Buffers and handles are cached per shape."""

    name = "ce_push"
    _cache = {}

    def __init__(self, k, v, rank, world, group, events=None):
        import torch.distributed._symmetric_memory as symm
        from cuda.bindings import driver as cu

        self.cu, self.rank, self.world, self.dev, self.events = cu, rank, world, k.device, events
        self.k, self.v = k, v
        kb, vb = k.numel() * k.element_size(), v.numel() * v.element_size()
        self.kb, self.vb = kb, vb
        slot = (kb + vb + 255) // 256 * 256
        gname = (group if group is not None else dist.group.WORLD).group_name
        key = (gname, slot, k.device.index)
        st = _PushExchange._cache.get(key)
        if st is None:
            buf = symm.empty(2 * world * slot, dtype=torch.uint8, device=k.device)  # [call parity][source rank]
            hdl = symm.rendezvous(buf, gname)
            peers = [hdl.get_buffer(p, (2 * world * slot,), torch.uint8) for p in range(world)]
            pads = [int(x) + 4096 for x in hdl.signal_pad_ptrs]  # second half of the pad: clear of torch's own barrier channels
            st = dict(buf=buf, hdl=hdl, peers=peers, pads=pads, epoch=0, stream=torch.cuda.Stream(device=k.device))
            hdl.barrier(channel=0)  # one-time: every rank's pads are zeroed and mapped before the first signal
            torch.cuda.synchronize(k.device)
            _PushExchange._cache[key] = st
        self.st, self.slot = st, slot
        st["epoch"] += 1
        self.epoch = st["epoch"]
        self.comm_stream = st["stream"]
        # signal pad words (uint32) on every rank: [src] = epoch of the data src pushed here; [world + dst] = epoch dst has consumed
        main = torch.cuda.current_stream(self.dev)
        self.comm_stream.wait_stream(main)  # K/V are ready
        with torch.cuda.stream(self.comm_stream):
            for t in range(1, world):
                dst = (rank + t) % world
                e0 = _mark(events, self.comm_stream)
                self._wait(self.comm_stream, st["pads"][rank] + 4 * (world + dst), self.epoch - 2)
                peer = st["peers"][dst]
                off = ((self.epoch & 1) * world + rank) * slot
                peer[off:off + kb].copy_(k.reshape(-1).view(torch.uint8), non_blocking=True)
                peer[off + kb:off + kb + vb].copy_(v.reshape(-1).view(torch.uint8), non_blocking=True)
                self._write(self.comm_stream, st["pads"][dst] + 4 * rank, self.epoch)
                e1 = _mark(events, self.comm_stream)
                if events is not None:
                    events.append(("comm", t - 1, e0, e1))  # slot t-1: the data hop t of the receiver needs

    def _check(self, res):
        err = res[0] if isinstance(res, tuple) else res
        if int(err) != 0:
            raise RuntimeError(f"CUDA driver call failed in _PushExchange: {err}")

    def _wait(self, stream, addr, value):
        self._check(self.cu.cuStreamWaitValue32(stream.cuda_stream, addr, value & 0xFFFFFFFF, self.cu.CUstreamWaitValue_flags.CU_STREAM_WAIT_VALUE_GEQ))

    def _write(self, stream, addr, value):
        self._check(self.cu.cuStreamWriteValue32(stream.cuda_stream, addr, value & 0xFFFFFFFF, 0))

    def begin(self, t):
        if t == 0:
            return self.k, self.v
        src = (self.rank - t) % self.world
        main = torch.cuda.current_stream(self.dev)
        self._wait(main, self.st["pads"][self.rank] + 4 * src, self.epoch)
        mine = self.st["buf"]
        off = ((self.epoch & 1) * self.world + src) * self.slot
        kk = mine[off:off + self.kb].view(self.k.dtype).view(self.k.shape)
        vv = mine[off + self.kb:off + self.kb + self.vb].view(self.v.dtype).view(self.v.shape)
        return kk, vv

    def end(self, t):
        pass

    def finish(self):
        # tell every source that its slot here may be overwritten by the next call; the caller's K/V stay in use by the
        # side stream until its pushes are done
        main = torch.cuda.current_stream(self.dev)
        for t in range(1, self.world):
            src = (self.rank - t) % self.world
            self._write(main, self.st["pads"][src] + 4 * (self.world + self.rank), self.epoch)
        main.wait_stream(self.comm_stream)


class _LocalExchange:
    """All shards live in this process (single-device simulation of the ring, used by tests and for debugging)."""

    def __init__(self, k_shards, v_shards, rank):
        self.k, self.v, self.rank = k_shards, v_shards, rank

    def begin(self, t):
        s = (self.rank - t) % len(self.k)
        return self.k[s], self.v[s]

    def end(self, t):
        pass

    def finish(self):
        pass


# --------------------------------------------------------------------------- the ring
TRANSPORTS = ("nccl", "ce_push")


def ring_flash_attn_func(q, k, v, softmax_scale=None, causal=True, group=None, backend=None, return_lse=False, _exchange=None,
                         _rank_world=None, transport="nccl", sm_margin=0, _events=None):
    """q, k, v: this rank's zig-zag shard, (b, 2c, h, d) / (b, 2c, h_k, d) / (b, 2c, h_k, dv), contiguous.
    Returns this rank's shard of the output (b, 2c, h, dv) [and LSE (b, h, 2c)].
    transport: "nccl" = K/V ring over NCCL send/recv (needs SMs for its copy kernels: `sm_margin` SMs are left free for
    them); "ce_push" = direct pushes into every peer's memory by the copy engines (no SMs, CUDA only).
    _events: a list that receives ("attn" | "comm", hop, start_event, end_event) for per-hop timing."""
    assert causal, "ring attention is built for the causal C5 case (non-causal needs no zig-zag; use flash_attn_func per hop)"
    if _rank_world is not None:
        rank, world = _rank_world
    else:
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    be = backend if backend is not None else Fa100Backend(softmax_scale, sm_margin if (world > 1 and transport == "nccl") else 0)
    b, rows2, h, d = q.shape
    c = rows2 // 2
    dv = v.shape[-1]
    assert rows2 % 2 == 0 and k.shape[1] == rows2 and v.shape[1] == rows2
    out = torch.empty((b, rows2, h, dv), dtype=q.dtype if q.dtype != torch.float8_e4m3fn else torch.bfloat16, device=q.device)
    out_lse = torch.empty((b, h, rows2), dtype=torch.float32, device=q.device)
    groups = {"lo": slice(0, c), "hi": slice(c, rows2)}
    acc = {g: be.alloc(b, h, c, dv, q.device) for g in groups}
    filled = {g: False for g in groups}
    plan = hop_plan(rank, world)
    last_use = {g: max(t for t, (_, work) in enumerate(plan) if any(w[0] == g for w in work)) for g in groups}
    if _exchange is not None:
        ex = _exchange
    elif transport == "ce_push" and world > 1 and k.is_cuda:
        ex = _PushExchange(k, v, rank, world, group, _events)
    else:
        assert transport in TRANSPORTS, transport
        ex = _RingExchange(k, v, rank, world, group, _events)
    timed = _events is not None and q.is_cuda

    for t, (src, work) in enumerate(plan):
        k_t, v_t = ex.begin(t)
        a0 = _mark(_events, torch.cuda.current_stream(q.device)) if timed else None
        for g, keys, is_causal in work:
            o, lse = acc[g]
            kk = k_t[:, :c] if keys == "lo" else k_t
            vv = v_t[:, :c] if keys == "lo" else v_t
            final = t == last_use[g]
            if filled[g] and getattr(be, "fused_merge", False):
                if final:
                    be.partial_merge(q[:, groups[g]], kk, vv, is_causal, o[0], lse[0], out[:, groups[g]], out_lse[:, :, groups[g]])
                else:
                    be.partial_merge(q[:, groups[g]], kk, vv, is_causal, o[0], lse[0])
                continue
            slot = 1 if filled[g] else 0
            be.partial(q[:, groups[g]], kk, vv, is_causal, o[slot], lse[slot])
            if filled[g]:
                if final:
                    be.merge(o, lse, out[:, groups[g]], out_lse[:, :, groups[g]])
                else:
                    be.merge(o, lse)
            elif final:
                be.finish_single(o, lse, out[:, groups[g]], out_lse[:, :, groups[g]])
            filled[g] = True
        if timed:
            _events.append(("attn", t, a0, _mark(_events, torch.cuda.current_stream(q.device))))
        ex.end(t)
    ex.finish()
    return (out, out_lse) if return_lse else out


def simulate_ring(q, k, v, world, softmax_scale=None, backend=None):
    """Run every rank of a `world`-way ring in this process on the full (b, s, h, d) tensors and reassemble the result:
    same schedule, partial launches and merges as the distributed path, no communication."""
    qs = [zigzag_shard(q, r, world) for r in range(world)]
    ks = [zigzag_shard(k, r, world) for r in range(world)]
    vs = [zigzag_shard(v, r, world) for r in range(world)]
    outs, lses = [], []
    for r in range(world):
        o, l = ring_flash_attn_func(qs[r], ks[r], vs[r], softmax_scale, True, None, backend, True, _LocalExchange(ks, vs, r), (r, world))
        outs.append(o)
        lses.append(l)
    return zigzag_unshard(outs, 1), zigzag_unshard(lses, 2)
