#!/bin/bash
# One --set full capture of the forward kernel on one workload: tools/gpu_ncu_one.sh TAG WORKLOAD [ENV=VAL ...]
set -u
mkdir -p gpurun_out
TAG=$1; WL=$2; shift 2
for kv in "$@"; do export "$kv"; done
CMD="python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras"
timeout 200 $CMD > gpurun_out/plain_${WL}_$TAG.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fa100_fwd -s 3 -c 1 -f -o gpurun_out/prof_${WL}_$TAG $CMD > gpurun_out/ncu_full_${WL}_$TAG.log 2>&1; echo "ncu full $WL rc=$?"
tail -2 gpurun_out/plain_${WL}_$TAG.log | cut -c1-300
