#!/bin/bash
# Multi-GPU box visit (gpurun --gpus N): ring tests over real NCCL / peer memory, the two-devices-one-process test, then
# the bench line at N ranks (C2 replicas + ring_c5 with both transports).  Every step under its own timeout.
set -u
mkdir -p gpurun_out
N=${1:-2}; TAG=${2:-r02_n$N}
nvidia-smi topo -m > gpurun_out/topo_$TAG.txt 2>&1
timeout 900 python -m pytest tests/test_ring_gpu.py tests/test_parity_matrix_gpu.py::test_two_devices_one_process -m gpu -q -x > gpurun_out/pytest_ring_$TAG.log 2>&1; echo "pytest ring rc=$?"; tail -15 gpurun_out/pytest_ring_$TAG.log
NCCL_DEBUG=WARN timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -c 3500 gpurun_out/bench_$TAG.json; tail -5 gpurun_out/bench_$TAG.err
