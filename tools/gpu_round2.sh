#!/bin/bash
# One single-GPU box visit of round 2: the no-alloc test alone, the parity suite, A/B of kernel variants, bench line with
# extras, library comparison (incl. the reference's CuTe-DSL kernel from baseline/_ref).
set -u
mkdir -p gpurun_out
TAG=${1:-r02a}; shift || true
LIBS=${*:-"libfa100.so libfa100.so:FA100_FOUR_ROLE=1"}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_$TAG.csv 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_alone_$TAG.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/pytest_noalloc_alone_$TAG.log
timeout 1200 python -m pytest tests -m gpu -q -x --maxfail=1 > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_$TAG.log
AB_ROUNDS=3 timeout 900 python tools/ab_time.py $LIBS > gpurun_out/ab_$TAG.log 2>&1; echo "ab rc=$?"; cat gpurun_out/ab_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err
timeout 900 python tools/compare_libs.py > gpurun_out/compare_$TAG.json 2> gpurun_out/compare_$TAG.err; echo "compare rc=$?"; head -c 1500 gpurun_out/compare_$TAG.json
