#!/bin/bash
# Ring (C5) at N ranks for several sm_margin values of the NCCL transport (the copy-engine transport is timed in every run).
set -u
mkdir -p gpurun_out
N=${1:-4}; TAG=${2:-r02_ring_n$N}; shift; shift
for M in ${*:-8 24}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29578 bench.py --gpus $N --workload c5 --steps 10 --warmup 3 --ring-sm-margin $M > gpurun_out/bench_${TAG}_m$M.json 2> gpurun_out/bench_${TAG}_m$M.err; echo "margin $M rc=$?"
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_${TAG}_m$M.json") if l.startswith("{")][-1])["ring_c5"]
print("n1 ms", round(d["n1_ms_per_step"],2), d.get("n1_clocks"))
for k,v in d["transports"].items():
    print(k, "sm_margin $M" if k=="nccl" else "", round(v["value"],1), "TF/s", round(v["ms_per_step"],3), "ms eff", round(v["strong_efficiency_vs_n1"],3), v["per_hop"]["attn_ms"], v["per_hop"]["transfer_ms"], "clocks", v.get("clocks"), "instr", v["per_hop"]["instrumented_step_ms"], "barriered", v["per_hop"].get("barriered_step_ms"), "exposed", v["per_hop"]["exposed_comm_ms"])
PY
done
