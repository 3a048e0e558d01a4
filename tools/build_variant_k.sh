#!/bin/bash
# Build a kernels.cu variant of the library (development aid): tools/build_variant_k.sh NAME [-DCANDF_...=...]...
# -> danbing-tk_b200/build/var/lib_NAME.so, timed on the GPU box by tools/exp_variants.sh NAME...
set -e
cd "$(dirname "$0")/../danbing-tk_b200"
name=$1; shift
mkdir -p build/var
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -cudart static "$@" -Xptxas -v -c csrc/kernels.cu -o build/var/kernels_$name.o 2> build/var/kernels_$name.ptxas
grep -A2 "candidate_fast_kernelILi2ELi5" build/var/kernels_$name.ptxas | grep -E "spill|registers" | head -2
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o build/var/lib_$name.so build/var/kernels_$name.o build/fastx.o build/bubbles.o build/thread.o build/dbt_abi.o build/noalloc.o build/pred.o build/host_pack.o -ldl -lpthread
