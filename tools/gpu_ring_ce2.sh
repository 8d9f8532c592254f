#!/bin/bash
# NCCL ring with NCCL_P2P_USE_CUDA_MEMCPY=1 and small SM margins: tools/gpu_ring_ce2.sh N
set -u
mkdir -p gpurun_out
N=${1:-2}
export NCCL_P2P_USE_CUDA_MEMCPY=1
for M in 2 0; do
  NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29582 \
    bench.py --gpus $N --workload c5 --steps 6 --warmup 3 --no-cpu-baseline --no-e2e --ring-sm-margin $M > gpurun_out/ring_ce2_n${N}_m$M.json 2> gpurun_out/ring_ce2_n${N}_m$M.err
  echo "margin $M rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ring_ce2_n${N}_m$M.json").read().strip().splitlines()[-1])
    for name, t in d["ring_c5"].get("transports", {}).items():
        print("   ", name, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in t.items() if not isinstance(v, (list, dict))})
        ph = t.get("per_hop")
        if ph: print("      per_hop", {k: [round(x, 2) for x in v] if isinstance(v, list) else v for k, v in ph.items()})
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/ring_ce2_n${N}_m$M.err").read()[-800:])
PY
done
