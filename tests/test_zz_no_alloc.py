"""GPU test of dbt_pack_reads_device_noalloc (csrc/noalloc.cu): the allocation-free variant of the read packer, which runs
pack_kernel inside a fixed file-scope device array instead of cudaMalloc'ed buffers.  It has no counterpart in the reference;
its oracle is the library's existing entry point for the same job, dbt_pack_reads (the host
packer, itself pinned to a numpy model in tests/test_abi.py).  Bit-exact, on two calls in a row -- the second one smaller and with
other reads, so anything the shared scratch kept from the first call would show.

Collected last on purpose: a failure here must not hide the other tests."""
import ctypes as C
import random

import numpy as np
import pytest

from oracle_abi import abi, rpgg_io

pytestmark = pytest.mark.gpu

E_CAPACITY = -4


def make_reads(rng, npairs, max_len, dirty):
    seqs = []
    for i in range(2 * npairs):
        n = max_len if rng.random() < 0.7 else rng.randrange(0, max_len + 1)
        s = [rng.choice("ACGT") for _ in range(n)]
        if dirty and rng.random() < 0.3:
            for _ in range(rng.randrange(1, 6)):
                if n:
                    s[rng.randrange(n)] = rng.choice("Nacgt@\n")
        seqs.append("".join(s))
    return rpgg_io.seqs_to_batch(seqs)


def pack_both(L, k, max_len, npairs, bases, off):
    nb = L.dbt_packed_bytes(max_len, npairs)
    exp = np.full(nb, 0xA5, np.uint8)
    got = np.full(nb, 0x5A, np.uint8)
    abi.check(L, L.dbt_pack_reads(k, max_len, npairs, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(exp, abi.u8p), 2), "dbt_pack_reads")
    abi.check(L, L.dbt_pack_reads_device_noalloc(0, k, max_len, npairs, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(got, abi.u8p)),
              "dbt_pack_reads_device_noalloc")
    return exp, got


def test_noalloc_packer_equals_the_existing_packer_on_two_calls_in_a_row():
    L = abi.load_library()
    if L.dbt_device_count() < 1:
        pytest.fail("no CUDA device: libdanbing_b200 has no CPU fallback")
    rng = random.Random(2024)
    # call 1: a batch near the limit, clean and dirty reads of every length
    b1, o1 = make_reads(rng, 12000, 150, dirty=True)
    assert L.dbt_noalloc_scratch_needed(150, 12000, int(o1[-1])) <= abi_const("DBT_NOALLOC_SCRATCH_BYTES")
    exp1, got1 = pack_both(L, 21, 150, 12000, b1, o1)
    assert got1.tobytes() == exp1.tobytes()
    # call 2, straight after: fewer pairs, another k and read length, offsets that do not start at 0
    b2, o2 = make_reads(rng, 777, 100, dirty=True)
    pad = 37
    b2 = np.concatenate([np.frombuffer(b"x" * pad, np.uint8), b2])
    o2 = o2 + np.uint64(pad)
    exp2, got2 = pack_both(L, 17, 100, 777, b2, o2)
    assert got2.tobytes() == exp2.tobytes()
    # and the first batch again: same bytes as before (nothing of call 2 leaks into it)
    exp3, got3 = pack_both(L, 21, 150, 12000, b1, o1)
    assert got3.tobytes() == exp1.tobytes() == exp3.tobytes()
    assert len(set(got1.tobytes())) > 100           # the output is not trivially constant


def abi_const(name):
    import os
    import re
    from oracle_abi import ROOT
    hdr = open(os.path.join(ROOT, "include", "danbing_b200.h")).read()
    m = re.search(r"#define\s+" + name + r"\s+\(?(\d+)u?(?:\s*<<\s*(\d+))?\)?", hdr)
    return int(m.group(1)) << int(m.group(2) or 0)


def test_noalloc_packer_rejects_what_does_not_fit_and_stays_usable():
    L = abi.load_library()
    rng = random.Random(7)
    max_pairs = abi_const("DBT_NOALLOC_MAX_PAIRS")
    assert max_pairs == 16384 and abi_const("DBT_NOALLOC_SCRATCH_BYTES") == 8 << 20
    # more pairs than the stated maximum
    n = max_pairs + 1
    bases, off = rpgg_io.seqs_to_batch(["ACGT" * 6] * (2 * n))
    out = np.zeros(L.dbt_packed_bytes(150, n), np.uint8)
    rc = L.dbt_pack_reads_device_noalloc(0, 21, 150, n, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p))
    assert rc == E_CAPACITY and b"at most" in L.dbt_last_error()
    # few enough pairs, but more bytes than the array holds (long reads)
    n = 12000
    bases, off = make_reads(rng, n, 400, dirty=False)
    out = np.zeros(L.dbt_packed_bytes(400, n), np.uint8)
    assert L.dbt_noalloc_scratch_needed(400, n, int(off[-1])) > abi_const("DBT_NOALLOC_SCRATCH_BYTES")
    rc = L.dbt_pack_reads_device_noalloc(0, 21, 400, n, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p))
    assert rc == E_CAPACITY and b"bytes of scratch" in L.dbt_last_error()
    # a read longer than max_read_len: reported, not packed
    bases, off = rpgg_io.seqs_to_batch(["ACGT" * 50, "ACGT" * 10])
    out = np.zeros(L.dbt_packed_bytes(150, 1), np.uint8)
    rc = L.dbt_pack_reads_device_noalloc(0, 21, 150, 1, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p))
    assert rc == E_CAPACITY and b"longer than max_read_len" in L.dbt_last_error()
    # null arguments and an empty batch
    assert L.dbt_pack_reads_device_noalloc(0, 21, 150, 1, None, abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p)) != 0
    assert L.dbt_pack_reads_device_noalloc(0, 21, 150, 0, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p)) == 0
    # after the refusals (the error word of the scratch was set by the long read): the next good call is right
    b, o = make_reads(rng, 500, 150, dirty=True)
    exp, got = pack_both(L, 25, 150, 500, b, o)
    assert got.tobytes() == exp.tobytes()
