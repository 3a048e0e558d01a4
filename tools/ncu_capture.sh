#!/bin/bash
# ncu captures of the round (run under gpurun; every bench command below has exited 0 without ncu before):
#   $1.launches.csv        launch list of the kernel-only steps (gpu__time_duration per launch, cold caches, serialised)
#   $1.l2.ncu-rep          --set full of pack / subfilter / candidate_fast / thread at configs[1] (1000 loci, 10 M pairs)
#   $1.hbm.ncu-rep         --set full of subfilter / candidate_fast / thread at 20 000 loci (tables far beyond L2), 2 M pairs
#   $1.pred.ncu-rep        --set full of the two pred kernels
set -x
out=${1:-gpurun_out/ncu}
common="--mode thread --no-cpu --e2e-input ascii --file-pairs 0 --no-rrand --steps 2 --warmup 1"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out.launches.csv python bench.py $common --hbm-loci 0 --pred-samples 0 > $out.launches.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:pack_kernel|subfilter_kernel|candidate_fast_kernel|thread_kernel' -c 4 -f -o $out.l2 python bench.py $common --hbm-loci 0 --pred-samples 0 > $out.l2.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:subfilter_kernel|candidate_fast_kernel|thread_kernel' -c 3 -f -o $out.hbm python bench.py $common --loci 20000 --pairs 2000000 --hbm-loci 0 --pred-samples 0 > $out.hbm.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:pred_' -c 3 -f -o $out.pred python bench.py $common --pairs 1000000 --hbm-loci 0 > $out.pred.log 2>&1
ls -la $out.*
