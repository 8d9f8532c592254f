#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native FlashAttention forward (driver contract).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c3nc|c4|c5] [--impl reference] [--no-extras]

The ONE JSON line is the headline workload (default C2).  So that every north_star config reaches the driver's records,
the default run attaches the others to the same line: at N = 1 `extra: {c3, c3nc, c4}` (FP8 causal / non-causal against
an FP8 peak measured in the same run, paged split-KV decode against the HBM peak), at N > 1 `ring_c5` (the 131072-token
context-parallel prefill over the N ranks: strong scaling, per-hop attention / transfer times, both transports).

Default workload = BASELINE.json configs[1] (C2): bf16, batch 8, 32 Q / 8 KV heads, seqlen 8192, hdim 128, causal.
A step is one forward over that batch.  `value` times the kernel with inputs resident in HBM (CUDA events on the
launching stream); `e2e` times the public flash_attn_func call with HOST (pinned) q/k/v copied in and the output
copied back every step.  `roofline` relates the kernel to the measured tensor peak (MEASURED_PEAKS.json);
`cpu_baseline` is the CPU oracle (a restatement of the reference's attention_ref) on a bounded sample.
With N > 1 (torchrun) every rank runs the same per-GPU workload (batch x heads shard with no communication:
weak scaling) and `value` is the aggregate.  `--impl reference` times the reference's CPU path (oracle port).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200"), os.path.join(ROOT, "oracle")]

import torch  # noqa: E402

WORKLOADS = {
    # name: (dtype, b, sq, sk, h, hk, d, causal, metric, unit, description)
    "c2": dict(dtype="bf16", b=8, sq=8192, sk=8192, h=32, hk=8, d=128, causal=True,
               desc="FA fwd bf16 Llama-3-8B shape: b8 32q/8kv s8192 d128 causal"),
    "c3": dict(dtype="e4m3", b=4, sq=16384, sk=16384, h=16, hk=16, d=128, causal=True,
               desc="FA fwd FP8 e4m3 per-tensor descale: b4 h16 s16384 d128 causal"),
    "c3nc": dict(dtype="e4m3", b=4, sq=16384, sk=16384, h=16, hk=16, d=128, causal=False,
                 desc="FA fwd FP8 e4m3 per-tensor descale: b4 h16 s16384 d128 non-causal"),
    "c4": dict(dtype="bf16", b=64, sq=1, sk=32768, h=32, hk=8, d=128, causal=False, page=256,
               desc="split-KV decode, paged KV (page 256): b64 32q/8kv sq1 sk32768 d128"),
    # one 131072-token sequence sharded over the ranks (zig-zag ring / context parallel): total work is fixed -> strong scaling
    "c5": dict(dtype="bf16", b=1, sq=131072, sk=131072, h=32, hk=32, d=128, causal=True, ring=True,
               desc="ring/context-parallel FA fwd bf16: 131072 tokens, 32 heads (MHA), d128, causal, zig-zag over the ranks"),
}


def flops(w):
    avg_k = (max(0, w["sk"] - w["sq"]) + w["sk"]) / 2 if w["causal"] else w["sk"]
    return w["b"] * w["h"] * 2 * w["sq"] * avg_k * (w["d"] + w["d"])


def decode_bytes(w):
    elt = 2
    kv = w["b"] * w["sk"] * w["hk"] * (w["d"] + w["d"]) * elt
    qo = 2 * w["b"] * w["sq"] * w["h"] * w["d"] * elt
    return kv + qo + w["b"] * (w["sk"] // w.get("page", 256)) * 4


def ncu_traffic(workload):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(workload)
    except Exception:  # noqa: BLE001
        return None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        return dict(hbm=m["hbm_gbs"], tensor=m["bf16_tflops"], tensor_sustained=m.get("bf16_tflops_sustained"), source="measured")
    except Exception:  # noqa: BLE001
        return dict(hbm=6650.0, tensor=1590.0, tensor_sustained=1400.0, source="fallback")


class ClockSampler:
    """SM clock and throttle reasons sampled through NVML every 5 ms while the timed region runs
    (B200_PROFILING.md clocks line; nvidia-smi -lms cannot sample an 80 ms region more than once)."""

    def __init__(self, index):
        self.index, self.rows, self.h, self.stop_flag, self.thread, self.nv = index, [], None, False, None, None

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            # NVML enumerates physical devices: map through CUDA_VISIBLE_DEVICES when it holds indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [x for x in vis.split(",") if x.strip().isdigit()]
            phys = int(ids[self.index]) if self.index < len(ids) else self.index
            self.h, self.nv = nv.nvmlDeviceGetHandleByIndex(phys), nv
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception as e:  # noqa: BLE001
            self.h, self.err = None, repr(e)

    def _pump(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.rows.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                  nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.005)

    def stop(self):
        if self.h is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvml unavailable: " + getattr(self, "err", "?")], samples=0)
        self.stop_flag = True
        self.thread.join(timeout=1.0)
        nv = self.nv
        sm = sorted(r[0] for r in self.rows)
        bits = 0
        for r in self.rows:
            bits |= r[1]
        names = [("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap)]
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=self.max_mhz,
                    reasons=[n for n, m in names if bits & m], samples=len(sm))


def make_inputs(w, device, seed, world=1):
    g = torch.Generator(device="cpu").manual_seed(seed)
    dt = torch.bfloat16
    b, sq, sk, h, hk, d = w["b"], w["sq"], w["sk"], w["h"], w["hk"], w["d"]
    t = dict()
    if w.get("ring"):  # this rank's zig-zag shard (2 chunks of s / 2N rows); synthetic, so it is drawn directly
        rows = sq // world
        mk = lambda *s: torch.randn(*s, dtype=dt, device=device).cpu()  # noqa: E731
        t["q"], t["k"], t["v"] = mk(b, rows, h, d), mk(b, rows, hk, d), mk(b, rows, hk, d)
        return t
    if "page" in w:
        ps = w["page"]
        npages = b * sk // ps
        t["q"] = torch.randn(b, sq, h, d, generator=g).to(dt)
        t["k"] = torch.randn(npages, ps, hk, d, dtype=torch.bfloat16, device=device)
        t["v"] = torch.randn(npages, ps, hk, d, dtype=torch.bfloat16, device=device)
        t["page_table"] = torch.randperm(npages, generator=g).to(torch.int32).view(b, sk // ps).to(device)
        t["cache_seqlens"] = torch.full((b,), sk, dtype=torch.int32, device=device)
        return t
    if w["dtype"] == "e4m3":
        mk = lambda *s: torch.randn(*s, dtype=torch.bfloat16, device=device).to(torch.float8_e4m3fn).cpu()  # noqa: E731
        t["q"], t["k"], t["v"] = mk(b, sq, h, d), mk(b, sk, hk, d), mk(b, sk, hk, d)
        for nm, val in (("q_descale", 0.5), ("k_descale", 0.75), ("v_descale", 1.25)):
            t[nm] = torch.full((b, hk), val, dtype=torch.float32, device=device)
    else:
        mk = lambda *s: torch.randn(*s, dtype=dt, device=device).cpu()  # noqa: E731
        t["q"], t["k"], t["v"] = mk(b, sq, h, d), mk(b, sk, hk, d), mk(b, sk, hk, d)
    return t


def cpu_sample(w, threads, slices, min_seconds=0.0):
    """CPU oracle on (batch, kv-group) slices of the workload -- `slices` of them, then more until `min_seconds` have
    passed; returns (seconds, flops done, slices done)."""
    import attention_oracle as A

    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    G = w["h"] // w["hk"]
    sq, sk = min(w["sq"], 8192), min(w["sk"], 8192)
    q = torch.randn(1, sq, G, w["d"], generator=g).to(torch.bfloat16)
    k = torch.randn(1, sk, 1, w["d"], generator=g).to(torch.bfloat16)
    v = torch.randn(1, sk, 1, w["d"], generator=g).to(torch.bfloat16)
    done, t0 = 0, time.perf_counter()
    while done < slices or time.perf_counter() - t0 < min_seconds:
        A.attention_oracle_sliced(q, k, v, rows_per_call=1024, causal=w["causal"])
        done += 1
    dt = time.perf_counter() - t0
    return dt, done * A.flops(1, G, sq, sk, w["d"], w["d"], w["causal"]), done


def run_reference(args, w, rank):
    """--impl reference: the reference's CPU path (oracle port of attention_ref) on all host threads."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    slices = 2
    for _ in range(min(args.warmup, 1)):
        cpu_sample(w, threads, 1)
    tot_t, tot_f = 0.0, 0.0
    for _ in range(args.steps):
        dt, fl, _ = cpu_sample(w, threads, slices, min_seconds=8.0)
        tot_t += dt
        tot_f += fl
    val = tot_f / tot_t / 1e12
    sample = f">= 8 s of (batch, kv-group) slices of {w['h'] // w['hk']} q-heads per step, seq capped at 8192, q-row-blocked"
    line = dict(impl="reference", metric=metric_name(w), value=val, unit="TFLOP/s", n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=tot_t / args.steps * 1e3, higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f32", data="synthetic", config=dict(workload=w["desc"]),
                cpu_baseline=dict(value=val, unit="TFLOP/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=val, unit="TFLOP/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


def metric_name(w):
    if "page" in w:
        return "decode HBM GB/s (split-KV paged decode)"
    if w.get("ring"):
        return "ring/context-parallel fwd TFLOPS (aggregate) bf16 hdim128 causal 131072 tokens"
    return f"fwd TFLOPS/GPU {w['dtype']} hdim128 {'causal' if w['causal'] else 'non-causal'}"




def fp8_peak(dev):
    """Dense e4m3 tensor peak of THIS box, measured the way MEASURED_PEAKS.json measures bf16: torch._scaled_mm 8192^3
    (cuBLASLt), best of 10 after warm-up (burst).  None when the op is not available."""
    try:
        n = 8192
        a = torch.randn(n, n, device=dev, dtype=torch.bfloat16).to(torch.float8_e4m3fn)
        b = torch.randn(n, n, device=dev, dtype=torch.bfloat16).to(torch.float8_e4m3fn).t()  # column-major B
        one = torch.ones((), device=dev, dtype=torch.float32)
        run = lambda: torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)  # noqa: E731
        for _ in range(3):
            run()
        best = None
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    except Exception:  # noqa: BLE001
        return None


class Ctx:
    """What every measurement needs: ranks, device, barrier."""

    def __init__(self, args):
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist

            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist

    def barrier(self):
        torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        t = torch.tensor(vals, dtype=torch.float64, device=self.dev)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()


def make_step(w, world, ring_kw=None):
    """The call a user makes for this workload (public API only)."""
    from flash_attn_interface import flash_attn_func, flash_attn_with_kvcache

    if w.get("ring"):
        import ring_attention as R

        kw = dict(ring_kw or {})
        return lambda t: R.ring_flash_attn_func(t["q"], t["k"], t["v"], causal=True, **kw)
    if "page" in w:
        return lambda t: flash_attn_with_kvcache(t["q"], t["k"], t["v"], cache_seqlens=t["cache_seqlens"],
                                                 page_table=t["page_table"], causal=False)

    def step(t):
        kw = {n: t[n] for n in ("q_descale", "k_descale", "v_descale") if n in t}
        return flash_attn_func(t["q"], t["k"], t["v"], causal=w["causal"], **kw)
    return step


def launches_per_step(w, world):
    if "page" in w:
        return 2  # forward (dual-stream partials) + combine
    if w.get("ring"):  # one attention launch per (hop, row group) of this rank's plan; merges are fused into them
        import ring_attention as R

        return sum(len(work) for _, work in R.hop_plan(0, world))
    return 1


def measure(ctx, w, steps, warmup, do_e2e, ring_kw=None, want_clocks=True):
    """Device-timed `steps` back-to-back steps (CUDA events on the launching stream, barrier + synchronize on both sides,
    max over ranks), then the end-to-end variant with HOST inputs / outputs.  Returns a dict of raw numbers."""
    import fa100_native as N

    rank, world, dev = ctx.rank, ctx.world, ctx.dev
    host = make_inputs(w, dev, seed=1234 + rank, world=world)
    pin = {k_: (v_.pin_memory() if v_.device.type == "cpu" else v_) for k_, v_ in host.items()}
    devt = {k_: v_.to(dev, non_blocking=True) for k_, v_ in pin.items()}
    torch.cuda.synchronize()
    step = make_step(w, world, ring_kw)
    for _ in range(warmup):
        out = step(devt)
    ctx.barrier()
    sampler = ClockSampler(ctx.local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.barrier()
    if rank == 0 and want_clocks:
        sampler.start()  # samples only while the timed steps run
    e0.record()
    for _ in range(steps):
        out = step(devt)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop() if (rank == 0 and want_clocks) else None
    ctx.barrier()
    ms = e0.elapsed_time(e1) / steps
    hang = N.hang_info()
    assert hang[0] == 0, f"kernel watchdog fired: {hang}"

    # e2e: through the public API with HOST inputs and a HOST result every step.  Each step copies its q/k/v from pinned
    # host memory, runs the forward and copies the output back; copy-in, compute and copy-out run on three streams with
    # double-buffered device inputs, so step i+1's upload and step i-1's download overlap step i's kernel (what a serving
    # loop over PCIe does).  Every byte of every step crosses the bus inside the timed region.
    e2e_ms, h2d, d2h = None, 0, 0
    if do_e2e:
        host_keys = [k_ for k_, v_ in pin.items() if v_.device.type == "cpu"]
        out_host = [torch.empty(out.shape, dtype=out.dtype).pin_memory() for _ in range(2)]
        h2d = sum(pin[k_].numel() * pin[k_].element_size() for k_ in host_keys)
        d2h = out_host[0].numel() * out_host[0].element_size()
        s_in, s_run, s_out = (torch.cuda.Stream(device=dev) for _ in range(3))
        dev_in = [{k_: torch.empty_like(devt[k_]) for k_ in host_keys} for _ in range(2)]
        in_ready = [torch.cuda.Event() for _ in range(2)]
        run_done = [torch.cuda.Event() for _ in range(2)]
        out_done = [torch.cuda.Event() for _ in range(2)]

        def e2e_steps(n):
            for i in range(n):
                sl = i % 2
                with torch.cuda.stream(s_in):
                    s_in.wait_event(run_done[sl])      # the forward that last read this input buffer has finished
                    for k_ in host_keys:
                        dev_in[sl][k_].copy_(pin[k_], non_blocking=True)
                    in_ready[sl].record(s_in)
                with torch.cuda.stream(s_run):
                    s_run.wait_event(in_ready[sl])
                    t = dict(devt)
                    t.update(dev_in[sl])
                    o = step(t)
                    run_done[sl].record(s_run)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(run_done[sl])
                    s_out.wait_event(out_done[sl])     # (host buffer reuse is ordered by the stream itself)
                    out_host[sl].copy_(o, non_blocking=True)
                    o.record_stream(s_out)
                    out_done[sl].record(s_out)
            s_run.wait_stream(s_in)
            s_run.wait_stream(s_out)

        e2e_steps(2)
        ctx.barrier()
        n_e2e = max(4, min(steps, 10))
        with torch.cuda.stream(s_run):
            e0.record(s_run)
            e2e_steps(n_e2e)
            e1.record(s_run)
        ctx.barrier()
        e2e_ms = e0.elapsed_time(e1) / n_e2e
    ms_max, e2e_max = ctx.max_over_ranks([ms, e2e_ms or 0.0])
    res = dict(ms=ms_max, e2e_ms=(e2e_max if e2e_ms is not None else None), h2d=h2d, d2h=d2h, clocks=clocks,
               launches=launches_per_step(w, world))
    if e2e_ms is not None:
        # per-rank bus rate of the e2e loop (copies of all ranks share the host's memory system and PCIe switches)
        gbs = ctx.max_over_ranks([-(h2d + d2h) / (e2e_ms * 1e-3) / 1e9])
        res["e2e_copy_gbs_slowest_rank"] = -gbs[0]
    del devt, out
    torch.cuda.empty_cache()
    return res


def measure_bwd(ctx, w, steps, warmup, peaks):
    """Backward of the same workload through torch.ops.flash_attn_3.bwd (three launches per step: D = rowsum(dO o O), the dQ
    kernel, the dK/dV kernel), counted as 2.5 x the forward's algorithmic FLOPs (hopper/benchmark_attn.py:62-73)."""
    import flash_attn_interface  # noqa: F401  (registers torch.ops.flash_attn_3)

    t = make_inputs(w, ctx.dev, seed=ctx.rank)
    q, k, v = (t[n].to(ctx.dev) for n in ("q", "k", "v"))
    out, lse, *_ = torch.ops.flash_attn_3.fwd(q, k, v, is_causal=w["causal"])
    g = torch.randn_like(out)
    step = lambda: torch.ops.flash_attn_3.bwd(g, q, k, v, out, lse, None, None, None, None, None, None, None, None, None, None, w["causal"])  # noqa: E731
    for _ in range(max(warmup, 3)):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = max(3, min(steps, 10))
    e0.record()
    for _ in range(n):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    tf = 2.5 * flops(w) / ms / 1e9
    peak = peaks["tensor"]
    return dict(metric="bwd TFLOPS/GPU bf16 hdim128 causal (2.5 x forward FLOPs)", value=tf, unit="TFLOP/s", ms_per_step=ms, dtype="bf16",
                gpu_launches=3 * n, workload=w["desc"] + ", backward (dq, dk, dv)",
                roofline=dict(bound="tensor", achieved=tf * 7 / 5, peak=peak, unit="TFLOP/s", frac=tf * 7 / 5 / peak, traffic=None,
                              note="achieved counts the 7 tile products the two kernels execute (S and dP are recomputed in both); value counts 5"))


def roofline_of(w, per_gpu, peaks, name, fp8_tflops=None):
    decode = "page" in w
    if decode:
        return dict(bound="hbm", achieved=per_gpu, peak=peaks["hbm"], unit="GB/s", frac=per_gpu / peaks["hbm"],
                    traffic=ncu_traffic(name), peak_source=peaks["source"] + " (copy)", kernel="fa100_fwd_kernel")
    fp8 = w["dtype"] == "e4m3"
    if fp8 and fp8_tflops:
        peak, src = fp8_tflops, "measured in this run: torch._scaled_mm e4m3 8192^3 best of 10 (burst)"
    elif fp8:
        peak, src = peaks["tensor"] * 2.0, peaks["source"] + " (bf16 burst x2: no fp8 GEMM on this box)"
    else:
        peak, src = peaks["tensor"], peaks["source"] + " (cuBLAS bf16 burst)"
    r = dict(bound="tensor", achieved=per_gpu, peak=peak, unit="TFLOP/s", frac=per_gpu / peak, traffic=ncu_traffic(name),
             peak_source=src, kernel="fa100_fwd_kernel")
    if peaks.get("tensor_sustained") and not fp8:
        r["frac_of_sustained"] = per_gpu / peaks["tensor_sustained"]
    if fp8:
        r["frac_of_2x_bf16_burst"] = per_gpu / (peaks["tensor"] * 2.0)
    return r


def measure_ring(ctx, steps, warmup):
    """C5 over the N ranks of this job: one 131072-token causal sequence, zig-zag sharded; strong scaling.  Both transports
    are timed (NCCL send/recv ring with SMs left free for its copy kernels; copy-engine pushes into peer memory), then one
    instrumented step per transport gives the per-hop attention / transfer times.  Every rank also runs the WHOLE problem
    alone (N = 1 arithmetic, no communication) so that the strong-scaling efficiency is measured inside one record."""
    import ring_attention as R

    w = WORKLOADS["c5"]
    world, rank, dev = ctx.world, ctx.rank, ctx.dev
    total = flops(w)
    out = dict(n_gpus=world, workload=w["desc"], unit="TFLOP/s", scaling="strong")
    # ---- N = 1 reference time of the same problem on one GPU (each rank, concurrently, no communication)
    g = torch.Generator(device="cpu").manual_seed(7)
    full = {n: torch.randn(1, w["sq"], w["h"] if n == "q" else w["hk"], w["d"], dtype=torch.bfloat16, device=dev) for n in ("q", "k", "v")}
    one = lambda: R.ring_flash_attn_func(full["q"], full["k"], full["v"], causal=True, _rank_world=(0, 1))  # noqa: E731
    for _ in range(2):
        one()
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
    e0.record()
    n1_steps = 3
    for _ in range(n1_steps):
        one()
    e1.record()
    torch.cuda.synchronize()
    if rank == 0:
        out["n1_clocks"] = sampler.stop()
    (t1,) = ctx.max_over_ranks([e0.elapsed_time(e1) / n1_steps])
    out["n1_ms_per_step"] = t1
    out["n1_tflops"] = total / (t1 * 1e-3) / 1e12
    del full
    torch.cuda.empty_cache()
    # ---- the ring, per transport
    rows = w["sq"] // world
    kv_bytes = 2 * w["b"] * rows * w["hk"] * w["d"] * 2
    results = {}
    ring_group = R.nccl_ring_group(ctx.args.ring_sm_margin) if not ctx.args.ring_default_comm else None
    for transport, kw in (("nccl", dict(transport="nccl", sm_margin=ctx.args.ring_sm_margin, group=ring_group)), ("ce_push", dict(transport="ce_push"))):
        try:
            m = measure(ctx, w, steps, warmup, do_e2e=False, ring_kw=kw, want_clocks=True)
            ms = m["ms"]
            r = dict(value=total / (ms * 1e-3) / 1e12, ms_per_step=ms, gpu_launches_per_step=m["launches"],
                     strong_efficiency_vs_n1=(t1 / world) / ms, clocks=m["clocks"])
            if transport == "nccl":
                r["sm_margin"] = ctx.args.ring_sm_margin
                r["nccl_comm"] = "default" if ring_group is None else f"dedicated: max_ctas={ctx.args.ring_sm_margin}, high-priority stream"
            # one instrumented step: events around every hop's attention launches and every transfer
            host = make_inputs(w, dev, seed=1234 + rank, world=world)
            t = {k_: v_.to(dev) for k_, v_ in host.items()}
            ev = []
            R.ring_flash_attn_func(t["q"], t["k"], t["v"], causal=True, **kw)  # warm: the instrumented step must not pay first-use costs
            ctx.barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            R.ring_flash_attn_func(t["q"], t["k"], t["v"], causal=True, _events=ev, **kw)
            s1.record()
            ctx.barrier()
            attn = [0.0] * world
            comm = [0.0] * world
            for kind, hop, a, b in ev:
                (attn if kind == "attn" else comm)[hop] += a.elapsed_time(b)
            (step_ms, not_attn) = ctx.max_over_ranks([s0.elapsed_time(s1), s0.elapsed_time(s1) - sum(attn)])
            # the same step repeated with a barrier in front of each: separates rank drift / coupling (gone with barriers)
            # from clock effects of a long run (still there)
            tot = 0.0
            for _ in range(5):
                ctx.barrier()
                s0.record()
                R.ring_flash_attn_func(t["q"], t["k"], t["v"], causal=True, **kw)
                s1.record()
                torch.cuda.synchronize()
                tot += s0.elapsed_time(s1)
            (barriered_ms,) = ctx.max_over_ranks([tot / 5])
            attn = ctx.max_over_ranks(attn)
            comm = ctx.max_over_ranks(comm)
            r["per_hop"] = dict(attn_ms=[round(x, 3) for x in attn], transfer_ms=[round(x, 3) for x in comm[:-1]],
                                bytes_per_hop=kv_bytes, instrumented_step_ms=round(step_ms, 3), barriered_step_ms=round(barriered_ms, 3),
                                exposed_comm_ms=round(not_attn, 3),
                                note="one instrumented step, max over ranks.  transfer_ms[t]: the K/V copy that hop t+1 needs (nccl: "
                                     "send/recv issued at hop t; ce_push: the push whose data the receiver uses at its hop t+1); "
                                     "exposed_comm_ms = that step's time - the sum of its hops' attention time (waits for K/V + launch gaps)")
            r["merge_launches"] = 0  # merges run in the attention epilogue (Fa100FwdParams.merge_partial)
            results[transport] = r
            del t
        except Exception as e:  # noqa: BLE001
            import traceback

            results[transport] = "unavailable: " + traceback.format_exc()[-900:]
            if transport == "nccl":
                raise
        torch.cuda.empty_cache()
    ok = {k_: v_ for k_, v_ in results.items() if isinstance(v_, dict)}
    best = max(ok, key=lambda k_: ok[k_]["value"])
    out.update(ok[best])
    out["transport"] = best
    out["transports"] = {k_: (dict(value=v_["value"], ms_per_step=v_["ms_per_step"], strong_efficiency_vs_n1=v_["strong_efficiency_vs_n1"],
                                   clocks=v_["clocks"], per_hop=v_["per_hop"], **({"sm_margin": v_["sm_margin"], "nccl_comm": v_["nccl_comm"]} if "nccl_comm" in v_ else {}))
                             if isinstance(v_, dict) else v_) for k_, v_ in results.items()}
    peaks = measured_peaks()
    out["roofline"] = dict(bound="tensor", achieved=out["value"] / world, peak=peaks["tensor"], unit="TFLOP/s per GPU",
                           frac=out["value"] / world / peaks["tensor"], peak_source=peaks["source"] + " (cuBLAS bf16 burst)",
                           nvlink_bytes_per_gpu_per_step=kv_bytes * (world - 1))
    return out


def line_for(ctx, name, m, peaks, fp8_tflops=None):
    """Fields of a bench line (or of an `extra` entry) for one measured workload."""
    w, world = WORKLOADS[name], ctx.world
    decode, ring = "page" in w, bool(w.get("ring"))
    ms, e2e_ms = m["ms"], m["e2e_ms"]
    if decode:
        per_gpu = decode_bytes(w) / (ms * 1e-3) / 1e9
        unit = "GB/s"
        e2e_val = decode_bytes(w) * world / (e2e_ms * 1e-3) / 1e9 if e2e_ms else None
    else:
        share = 1.0 / world if ring else 1.0  # ring: the ranks split ONE problem
        per_gpu = flops(w) * share / (ms * 1e-3) / 1e12
        unit = "TFLOP/s"
        e2e_val = flops(w) * share * world / (e2e_ms * 1e-3) / 1e12 if e2e_ms else None
    d = dict(metric=metric_name(w), value=per_gpu * world, unit=unit, ms_per_step=ms,
             dtype={"bf16": "bf16", "e4m3": "fp8_e4m3"}[w["dtype"]],
             config=dict(workload=w["desc"]),
             notes=dict(per_gpu=("1/N of the sequence per GPU, K/V exchanged between the ranks" if ring else
                                 "full workload on every GPU (batch x heads sharding, no communication)"),
                        l2="inputs (>= 0.67 GB) exceed the 126 MB L2; no flush between steps"),
             clocks=m["clocks"], gpu_launches=ctx.args.steps * m["launches"],
             e2e=dict(value=e2e_val, unit=unit, h2d_bytes_per_step=m["h2d"], d2h_bytes_per_step=m["d2h"],
                      copy_gbs_slowest_rank=m.get("e2e_copy_gbs_slowest_rank")),
             roofline=roofline_of(w, per_gpu, peaks, name, fp8_tflops))
    return d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="fa100", choices=["fa100", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="only the headline workload (no extra / ring_c5 attachments)")
    ap.add_argument("--ring-default-comm", action="store_true", help="NCCL ring over the default communicator instead of the max_ctas-capped one")
    ap.add_argument("--ring-sm-margin", type=int, default=16, help="SMs the attention kernel leaves to NCCL's send/recv kernels in the ring")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "fa100" else args.warmup
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w, int(os.environ.get("RANK", "0")))
        return

    import fa100_native  # noqa: F401  (fails loudly when libfa100.so is missing: there is no fallback)

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    ctx = Ctx(args)
    rank, world = ctx.rank, ctx.world
    peaks = measured_peaks()
    ring = bool(w.get("ring"))
    fp8_tflops = fp8_peak(ctx.dev) if (w["dtype"] == "e4m3" or (world == 1 and not args.no_extras)) else None

    if ring and world > 1:
        r = measure_ring(ctx, args.steps, args.warmup)
        if rank == 0:
            line = dict(metric=metric_name(w), value=r["value"], unit="TFLOP/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
                        ms_per_step=r["ms_per_step"], higher_is_better=True, scaling="strong", vs_baseline=None, dtype="bf16",
                        data="synthetic", config=dict(workload=w["desc"], transport=r["transport"]), clocks=r.get("clocks"),
                        gpu_launches=args.steps * r["gpu_launches_per_step"], e2e=None, roofline=r["roofline"], cpu_baseline=None, ring_c5=r)
            print(json.dumps(line), flush=True)
    else:
        m = measure(ctx, w, args.steps, args.warmup, do_e2e=not args.no_e2e)
        line = line_for(ctx, args.workload, m, peaks, fp8_tflops)
        line.update(n_gpus=world, steps=args.steps, warmup=args.warmup, higher_is_better=True, scaling="strong" if ring else "weak",
                    vs_baseline=None, data="synthetic")
        if fp8_tflops is not None:
            line["fp8_peak_tflops_measured"] = fp8_tflops
        if not args.no_extras and args.workload == "c2":
            if world == 1:
                # the other single-GPU north_star configs, each with its own roofline, on the same line
                extra = {}
                for name in ("c3", "c3nc", "c4"):
                    try:
                        mx = measure(ctx, WORKLOADS[name], args.steps, args.warmup, do_e2e=False, want_clocks=True)
                        ex = line_for(ctx, name, mx, peaks, fp8_tflops)
                        extra[name] = {k_: ex[k_] for k_ in ("metric", "value", "unit", "ms_per_step", "dtype", "clocks", "gpu_launches", "roofline")}
                        extra[name]["workload"] = WORKLOADS[name]["desc"]
                    except Exception as e:  # noqa: BLE001
                        extra[name] = "failed: " + repr(e)[:300]
                try:
                    extra["bwd"] = measure_bwd(ctx, w, args.steps, args.warmup, peaks)
                except Exception as e:  # noqa: BLE001
                    extra["bwd"] = "failed: " + repr(e)[:300]
                line["extra"] = extra
            else:
                line["ring_c5"] = measure_ring(ctx, max(3, min(args.steps, 10)), args.warmup)
        if rank == 0:
            if not args.no_cpu_baseline and world == 1:
                threads = os.cpu_count() or 1
                dt, fl, slices = cpu_sample(w, threads, 4, min_seconds=12.0)
                line["cpu_baseline"] = dict(value=fl / dt / 1e12, unit="TFLOP/s", cores=threads, kind="port",
                                            sample=f"{slices} (batch, kv-group) slices ({slices * (w['h'] // w['hk'])} of "
                                                   f"{w['b'] * w['h']} q-head rows), seq capped at 8192, {dt:.1f} s")
            else:
                line["cpu_baseline"] = None
            order = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                     "dtype", "data", "config", "clocks", "gpu_launches", "e2e", "roofline", "cpu_baseline"]
            line = {**{k_: line[k_] for k_ in order if k_ in line}, **{k_: v_ for k_, v_ in line.items() if k_ not in order}}
            print(json.dumps(line), flush=True)
    if ctx.dist is not None:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
