#!/bin/bash
# One GPU-box visit: parity tests, bench lines, ncu launch list + one full capture of the forward kernel.
set -u
mkdir -p gpurun_out
TAG=${1:-r1}
timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 -x > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_$TAG.log
for wl in c2 c3 c3nc c4; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 > gpurun_out/bench_${wl}_$TAG.json 2> gpurun_out/bench_${wl}_$TAG.err; echo "bench $wl rc=$?"; tail -c 1500 gpurun_out/bench_${wl}_$TAG.json; tail -3 gpurun_out/bench_${wl}_$TAG.err
done
CMD="python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_c2_$TAG.csv $CMD > gpurun_out/ncu1_$TAG.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fa100_fwd -s 3 -c 1 -f -o gpurun_out/prof_c2_$TAG $CMD > gpurun_out/ncu2_$TAG.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
