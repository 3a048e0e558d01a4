#!/bin/bash
# Development aid: a library built with -DDBT_DEBUG_DIRTY (every aln record's pad[0] says which mates the fused / fast threading
# handed on to thread_kernel) -> danbing-tk_b200/build/var/lib_dirty.so, read by tools/dirty_breakdown.py.  Not part of `make all`.
set -e
cd "$(dirname "$0")/../danbing-tk_b200"
mkdir -p build/var
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -cudart static -DDBT_DEBUG_DIRTY"
nvcc $F -c csrc/kernels.cu -o build/var/kernels_dirty.o &
nvcc $F -c csrc/thread.cu -o build/var/thread_dirty.o &
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o build/var/lib_dirty.so build/var/kernels_dirty.o build/fastx.o build/bubbles.o build/var/thread_dirty.o build/dbt_abi.o build/noalloc.o build/pred.o build/host_pack.o -ldl -lpthread
ls -la build/var/lib_dirty.so
