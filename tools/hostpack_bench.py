"""Host packer throughput on this box (development aid): dbt_pack_reads GB/s of ASCII in, by thread count."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
from oracle_abi import abi
L = abi.load_library()
P, RLEN = 2_500_000, 150
rng = np.random.default_rng(1)
host, own = abi.pinned_array(L, P * 2 * RLEN)
host[:] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, P * 2 * RLEN, dtype=np.uint8)]
off = (np.arange(2 * P + 1, dtype=np.uint64) * RLEN)
out, own2 = abi.pinned_array(L, L.dbt_packed_bytes(RLEN, P))
print("cpus", len(os.sched_getaffinity(0)))
for nt in (1, 2, 4, 8, 12, 16):
    best = 1e9
    for _ in range(3):
        t = time.perf_counter()
        rc = L.dbt_pack_reads(21, RLEN, P, abi.ptr(host, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p), nt)
        best = min(best, time.perf_counter() - t)
    print(f"threads {nt:2d}: {best*1e3:7.2f} ms  {P*2*RLEN/best/1e9:6.2f} GB/s ASCII in")
# plain memcpy of the same bytes for scale
dst = np.empty_like(host)
t = time.perf_counter(); dst[:] = host; dt = time.perf_counter() - t
print(f"numpy copy (1 thread): {P*2*RLEN/dt/1e9:.2f} GB/s")
