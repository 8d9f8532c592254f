#!/bin/bash
# Race hunting without compute-sanitizer: the parity tests that cover every hand-off pattern, on the jitter build
# (random pauses of up to 2 us in front of every barrier wait, arrive and tcgen05.commit; -DFA100_JITTER on all six .cu files:
# make -C flash-attention-annotated_b200 libfa100_jitter.so).
set -u
mkdir -p gpurun_out
FA100_LIB=libfa100_jitter.so timeout 1500 python -m pytest tests/test_fwd_gpu.py tests/test_ring_gpu.py tests/test_zz_no_alloc.py tests/test_bwd_gpu.py tests/test_parity_matrix_gpu.py -m gpu -q -x \
  -k "test_output or test_masks or test_varlen or test_kvcache or test_fp8 or test_two_cta or test_four_role or test_split_rows or test_bitwise or test_decode_many or test_ragged or test_cuda_graph or test_large_head or test_growing or simulated_ring or fused_merge or noalloc or bwd or dynamic_schedule or tma_store or test_qv" \
  > gpurun_out/pytest_jitter.log 2>&1
echo "jitter pytest rc=$?"; tail -4 gpurun_out/pytest_jitter.log
