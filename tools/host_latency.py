"""Host cost of one call through the public API (review item: "host layer cost is unmeasured").
For each decode shape: microseconds of HOST time to enqueue one flash_attn_with_kvcache call (no synchronisation; the GPU
queue never runs dry because the kernel is shorter than the enqueue only for the tiny shape), and the synchronous latency
(call + torch.cuda.synchronize), with and without a precomputed scheduler_metadata.  Prints one JSON object.
  python tools/host_latency.py > gpurun_out/host_latency.json"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "flash-attention-annotated_b200")]
import torch  # noqa: E402
from flash_attn_interface import flash_attn_func, flash_attn_with_kvcache, get_scheduler_metadata  # noqa: E402


def host_us(fn, n=300):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    return (t1 - t0) / n * 1e6, (t2 - t0) / n * 1e6


def sync_us(fn, n=100):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


def decode_case(b, sk, h, hk, d, ps):
    dev = "cuda"
    npages = b * sk // ps
    q = torch.randn(b, 1, h, d, device=dev, dtype=torch.bfloat16)
    kp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(npages, ps, hk, d, device=dev, dtype=torch.bfloat16)
    table = torch.randperm(npages, device=dev).to(torch.int32).view(b, sk // ps)
    lens = torch.full((b,), sk, dtype=torch.int32, device=dev)
    meta = get_scheduler_metadata(b, 1, sk, h, hk, d, lens, qkv_dtype=torch.bfloat16, page_size=ps)
    plain = lambda: flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens, page_table=table)  # noqa: E731
    with_meta = lambda: flash_attn_with_kvcache(q, kp, vp, cache_seqlens=lens, page_table=table, scheduler_metadata=meta)  # noqa: E731
    graph = torch.cuda.CUDAGraph()
    with_meta()
    torch.cuda.synchronize()
    with torch.cuda.graph(graph):
        with_meta()
    res = {}
    for name, fn in (("plain", plain), ("scheduler_metadata", with_meta), ("cuda_graph_replay", graph.replay)):
        enq, thr = host_us(fn)
        res[name] = dict(host_enqueue_us=round(enq, 1), us_per_call_back_to_back=round(thr, 1), sync_latency_us=round(sync_us(fn), 1))
    return res


def main():
    out = {"device": torch.cuda.get_device_name(0)}
    out["C4 decode b64 32q/8kv sk32768 page256"] = decode_case(64, 32768, 32, 8, 128, 256)
    out["decode b1 32q/8kv sk4096 page256"] = decode_case(1, 4096, 32, 8, 128, 256)
    q = torch.randn(1, 512, 8, 128, device="cuda", dtype=torch.bfloat16)
    enq, thr = host_us(lambda: flash_attn_func(q, q, q, causal=True))
    out["prefill b1 h8 s512 (flash_attn_func)"] = dict(host_enqueue_us=round(enq, 1), us_per_call_back_to_back=round(thr, 1))
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
