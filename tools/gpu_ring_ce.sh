#!/bin/bash
# NCCL send/recv ring with and without NCCL's copy-engine path for peer-to-peer messages (NCCL_P2P_USE_CUDA_MEMCPY=1):
# tools/gpu_ring_ce.sh N TAG  ->  gpurun_out/ring_ce_TAG_{kernel,memcpy}.json (bench.py --workload c5 lines)
set -u
mkdir -p gpurun_out
N=${1:-2}; TAG=${2:-n$N}
for MODE in kernel memcpy; do
  if [ $MODE = memcpy ]; then export NCCL_P2P_USE_CUDA_MEMCPY=1; else unset NCCL_P2P_USE_CUDA_MEMCPY; fi
  NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29581 \
    bench.py --gpus $N --workload c5 --steps 6 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ring_ce_${TAG}_$MODE.json 2> gpurun_out/ring_ce_${TAG}_$MODE.err
  echo "$MODE rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ring_ce_${TAG}_$MODE.json").read().strip().splitlines()[-1])
    r = d["ring_c5"]
    print("$MODE", {k: (round(v, 2) if isinstance(v, float) else v) for k, v in r.items() if k in ("value", "ms_per_step", "transport", "efficiency_vs_n1")})
    for name, t in r.get("transports", {}).items():
        print("   ", name, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in t.items() if not isinstance(v, (list, dict))})
except Exception as e:
    print("$MODE parse failed", e); print(open("gpurun_out/ring_ce_${TAG}_$MODE.err").read()[-800:])
PY
done
