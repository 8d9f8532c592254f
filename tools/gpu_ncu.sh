#!/bin/bash
# ncu evidence for profiles/: launch list of the bench command, then --set full captures of the forward kernel for C2, C3, C4.
# Every profiled command first runs plain (must exit 0), as the profiling guide requires.
set -u
mkdir -p gpurun_out
TAG=${1:-r02}
CMD="python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extras"
timeout 200 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_c2_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1; echo "ncu launch list rc=$?"
for WL in c2 c3 c4; do
  CMD="python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras"
  timeout 200 $CMD > gpurun_out/plain_${WL}_$TAG.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:fa100_fwd -s 3 -c 1 -f -o gpurun_out/prof_${WL}_$TAG $CMD > gpurun_out/ncu_full_${WL}_$TAG.log 2>&1; echo "ncu full $WL rc=$?"
done
ls -la gpurun_out | grep -E "ncu-rep|launches" 
