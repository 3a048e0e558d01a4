#!/bin/bash
# Kernel-variant timing on the GPU box (development aid): for each lib variant built into
# danbing-tk_b200/build/var/lib_<name>.so run the threading bench without the CPU arm and print the per-kernel times.
cd "$(dirname "$0")/.."
for v in "$@"; do
  DBT_LIB=$PWD/danbing-tk_b200/build/var/lib_$v.so python bench.py --mode thread --steps 3 --warmup 2 --no-cpu --e2e-input ascii --hbm-loci 0 --file-pairs 0 > gpurun_out/exp_$v.json 2> gpurun_out/exp_$v.err || { echo "$v FAILED"; tail -3 gpurun_out/exp_$v.err; continue; }
  python - "$v" <<'P'
import json,sys
v=sys.argv[1]; d=json.load(open(f'gpurun_out/exp_{v}.json')); r=d['roofline']
print(v, 'step_ms %.2f'%d['ms_per_step'], 'tfast_ms %.2f'%r['per_kernel_ms']['thread_fast'], 'thread_ms %.2f'%r['per_kernel_ms']['thread'], 'cand %.2f'%r['per_kernel_ms']['candidate_fast'], 'e2e_ms %.1f'%d['e2e']['ms_per_step'], 'ninc', d['counters_per_step_rank0']['ninc'])
P
done
