#!/bin/bash
# Build an experiment variant of the library: tools/build_variant.sh NAME "-DFA100_POLY_FP8=2 ..." -> libfa100_NAME.so
# (only fa100_fwd.cu is recompiled; the other objects come from the product build).  Time with tools/ab_time.py.
set -e
NAME=$1; FLAGS=${2:-}
cd "$(dirname "$0")/../flash-attention-annotated_b200"
make -j4 > /dev/null
mkdir -p build/var_$NAME
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --use_fast_math -Xcompiler -fPIC -Xptxas -v $FLAGS \
  -c csrc/fa100_fwd.cu -o build/var_$NAME/fa100_fwd.o 2> build/var_$NAME/ptxas.log || (cat build/var_$NAME/ptxas.log; exit 1)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o libfa100_$NAME.so build/fa100_api.o build/var_$NAME/fa100_fwd.o build/fa100_combine.o build/fa100_append.o build/fa100_noalloc.o build/fa100_bwd.o -cudart static
echo "built libfa100_$NAME.so"
