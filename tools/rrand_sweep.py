"""R_rand sweep: random aligned gathers of 16 / 32 / 64 / 128 bytes from a table far beyond L2 (dbt_random_gather_bench).
What one random access costs as a function of its size decides the bucket size of the index tables."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from oracle_abi import abi  # noqa: E402  (loader of the package by path; nothing of oracle/ is executed)

L = abi.load_library()
out = {}
for tb in (4 << 30, 16 << 30):
    for sb, coop in ((16, 0), (32, 0), (64, 0), (128, 0), (32, 1), (64, 1), (128, 1), (32, 2)):
        ms = C.c_float()
        n = 1 << 28
        abi.check(L, L.dbt_random_gather_bench(0, tb, sb | (0x10000 if coop == 1 else 0x20000 if coop == 2 else 0), n, 3, C.byref(ms)), "gather")
        out[f"table_{tb >> 30}GiB_{sb}B" + ("_one_request_per_chunk" if coop == 1 else "_one_256bit_load_per_thread" if coop == 2 else "_one_thread_per_chunk")] = {
            "Ggathers_per_s": n / ms.value / 1e6, "useful_GBps": n * sb / ms.value / 1e6}
print(json.dumps(out, indent=1))
