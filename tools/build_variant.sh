#!/bin/bash
# Build a thread_kernel variant of the library (development aid): tools/build_variant.sh NAME [-DTHR_...=...]...
# -> danbing-tk_b200/build/var/lib_NAME.so, timed on the GPU box by tools/exp_variants.sh NAME...
set -e
cd "$(dirname "$0")/../danbing-tk_b200"
name=$1; shift
mkdir -p build/var
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -cudart static "$@" -Xptxas -v -c csrc/thread.cu -o build/var/thread_$name.o 2> build/var/thread_$name.ptxas
grep -A2 "thread_kernelILi2ELi5" build/var/thread_$name.ptxas | grep -E "spill|registers" | head -2
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o build/var/lib_$name.so build/kernels.o build/fastx.o build/bubbles.o build/var/thread_$name.o build/dbt_abi.o build/noalloc.o build/pred.o build/host_pack.o -ldl -lpthread
