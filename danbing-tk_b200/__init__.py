"""danbing-tk_b200: B200-native `danbing-tk align` hot path.

The product is native: csrc/ (CUDA kernels + C ABI -> libdanbing_b200.so) and host/ (the drop-in
`danbing-tk` CLI, loaders, FASTA batcher, writers).  The Python modules here are plumbing for
tests/ and bench.py (ctypes mirror of include/danbing_b200.h, numpy file readers, workload synth).
"""
from . import abi, rpgg_io, synth  # noqa: F401
