import sys, os, tempfile
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import bench
from oracle_abi import abi
import thread_common as tc
wd = tempfile.mkdtemp(dir='/dev/shm')
ix, pref, pool, hap_off = bench.build_workload(200, 20261104, wd)
p = abi.default_params(k=21, cth=45, flags=abi.F_THREADING | abi.F_CORRECTION | abi.F_ALN_MIN, max_batch_pairs=4096)
p.thread_cth, p.max_ncorrection = 85, 3
ctx = abi.Context(p, abi.make_index_struct(ix))
print("bench workload merged:", ctx.L.dbt_graph_merged(ctx.h))
_, ixt, tseqs, _ = tc.thread_case(write_text=False, frac_bg=0.2, nloci=8, seed=7, npairs=600)
pt = tc.thread_params(extra_flags=abi.F_ALN_MIN, max_batch_pairs=4096)
ctx2 = abi.Context(pt, abi.make_index_struct(ixt))
print("test case merged:", ctx2.L.dbt_graph_merged(ctx2.h))
