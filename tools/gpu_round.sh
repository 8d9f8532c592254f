#!/bin/bash
# One GPU-box visit: parity tests, bench lines, then ONE ncu pass (launch list by default; `full <wl>` = --set full capture).
set -u
mkdir -p gpurun_out
TAG=${1:-r1}; MODE=${2:-list}; WL=${3:-c2}
if [ "$MODE" != "full" ]; then
  timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 -x > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_$TAG.log
  for wl in c2 c3 c3nc c4; do
    timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 > gpurun_out/bench_${wl}_$TAG.json 2> gpurun_out/bench_${wl}_$TAG.err; echo "bench $wl rc=$?"; tail -c 1800 gpurun_out/bench_${wl}_$TAG.json; tail -3 gpurun_out/bench_${wl}_$TAG.err
  done
  timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>&1; echo "bench ref rc=$?"; tail -c 600 gpurun_out/bench_ref_$TAG.json
  CMD="python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
  timeout 200 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_c2_$TAG.csv $CMD > gpurun_out/ncu1_$TAG.log 2>&1; echo "ncu launches rc=$?"
else
  CMD="python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
  timeout 200 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:fa100_fwd -s 3 -c 1 -f -o gpurun_out/prof_${WL}_$TAG $CMD > gpurun_out/ncu2_$TAG.log 2>&1; echo "ncu full rc=$?"
  tail -2 gpurun_out/plain_$TAG.log | cut -c1-400
fi
ls gpurun_out | tail -5
