#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native FlashAttention forward (driver contract).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c3nc|c4] [--impl reference]

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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="fa100", choices=["fa100", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "fa100" else args.warmup
    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, w, rank)
        return

    import fa100_native as N
    from flash_attn_interface import flash_attn_func, flash_attn_with_kvcache

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    host = make_inputs(w, dev, seed=1234 + rank, world=world)
    pin = {k_: (v_.pin_memory() if v_.device.type == "cpu" else v_) for k_, v_ in host.items()}
    devt = {k_: v_.to(dev, non_blocking=True) for k_, v_ in pin.items()}
    torch.cuda.synchronize()
    decode = "page" in w

    ring = bool(w.get("ring"))
    if ring:
        import ring_attention as R

    def step(t):
        if ring:
            return R.ring_flash_attn_func(t["q"], t["k"], t["v"], causal=True)
        if decode:
            return flash_attn_with_kvcache(t["q"], t["k"], t["v"], cache_seqlens=t["cache_seqlens"], page_table=t["page_table"],
                                           causal=False)
        kw = {n: t[n] for n in ("q_descale", "k_descale", "v_descale") if n in t}
        return flash_attn_func(t["q"], t["k"], t["v"], causal=w["causal"], **kw)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        out = step(devt)
    barrier()
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if rank == 0:
        sampler.start()  # samples only while the timed steps run
    e0.record()
    for _ in range(args.steps):
        out = step(devt)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    hang = N.hang_info()
    assert hang[0] == 0, f"kernel watchdog fired: {hang}"

    # e2e: through the public API with HOST inputs and a HOST result every step.  Each step copies its q/k/v from pinned
    # host memory, runs the forward and copies the output back; copy-in, compute and copy-out run on three streams with
    # double-buffered device inputs, so step i+1's upload and step i-1's download overlap step i's kernel (what a serving
    # loop over PCIe does).  Every byte of every step crosses the bus inside the timed region.
    e2e_ms, h2d, d2h = None, 0, 0
    if not args.no_e2e:
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
        barrier()
        n_e2e = max(4, min(args.steps, 10))
        with torch.cuda.stream(s_run):
            e0.record(s_run)
            e2e_steps(n_e2e)
            e1.record(s_run)
        barrier()
        e2e_ms = e0.elapsed_time(e1) / n_e2e
    vals = torch.tensor([ms, e2e_ms or 0.0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    ms, e2e_ms = vals[0].item(), (vals[1].item() if e2e_ms is not None else None)

    launches_per_step = 2 if decode else 1
    if ring:  # partial-attention launches of this rank's hop plan + one merge per partial after the first of a row group
        n_part = sum(len(work) for _, work in R.hop_plan(0, world))
        launches_per_step = n_part + max(0, n_part - 2)
    if rank == 0:
        peaks = measured_peaks()
        if decode:
            per_gpu = decode_bytes(w) / (ms * 1e-3) / 1e9
            unit, bound, peak = "GB/s", "hbm", peaks["hbm"]
            e2e_val = decode_bytes(w) * world / (e2e_ms * 1e-3) / 1e9 if e2e_ms else None
        else:
            share = 1.0 / world if ring else 1.0  # ring: the ranks split ONE problem
            per_gpu = flops(w) * share / (ms * 1e-3) / 1e12
            unit, bound, peak = "TFLOP/s", "tensor", peaks["tensor"] * (2.0 if w["dtype"] == "e4m3" else 1.0)
            e2e_val = flops(w) * share * world / (e2e_ms * 1e-3) / 1e12 if e2e_ms else None
        line = dict(metric=metric_name(w), value=per_gpu * world, unit=unit, n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms, higher_is_better=True, scaling="strong" if ring else "weak", vs_baseline=None,
                    dtype={"bf16": "bf16", "e4m3": "fp8_e4m3"}[w["dtype"]], data="synthetic",
                    config=dict(workload=w["desc"], per_gpu=("1/N of the sequence per GPU, K/V ring over NCCL send/recv" if ring else
                                                            "full workload on every GPU (batch x heads sharding, no communication)"),
                                l2="inputs (>= 0.67 GB) exceed the 126 MB L2; no flush between steps"),
                    clocks=clocks, gpu_launches=args.steps * launches_per_step,
                    e2e=dict(value=e2e_val, unit=unit, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h),
                    roofline=dict(bound=bound, achieved=per_gpu, peak=peak, unit=unit, frac=per_gpu / peak, traffic=ncu_traffic(args.workload),
                                  peak_source=peaks["source"] + (" (bf16 burst x2 for fp8)" if w["dtype"] == "e4m3" and not decode else ""),
                                  frac_of_sustained=(per_gpu / (peaks["tensor_sustained"] * (2.0 if w["dtype"] == "e4m3" else 1.0))
                                                     if bound == "tensor" and peaks["tensor_sustained"] else None),
                                  kernel="fa100_fwd_kernel"))
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            dt, fl, slices = cpu_sample(w, threads, 4, min_seconds=12.0)
            line["cpu_baseline"] = dict(value=fl / dt / 1e12, unit="TFLOP/s", cores=threads, kind="port",
                                        sample=f"{slices} (batch, kv-group) slices ({slices * (w['h'] // w['hk'])} of "
                                               f"{w['b'] * w['h']} q-head rows), seq capped at 8192, {dt:.1f} s")
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
