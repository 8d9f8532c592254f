#!/bin/bash
# ncu full capture of the forward kernel on workload $1 (default c2); report -> gpurun_out/prof_$1_$2.ncu-rep
WL=${1:-c2}; TAG=${2:-x}
CMD="python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 120 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 280 ncu --set full --clock-control none --import-source on -k regex:fa100_fwd -s 3 -c 1 -f -o gpurun_out/prof_${WL}_$TAG $CMD > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/plain_$TAG.log | cut -c1-300
