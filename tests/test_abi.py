"""CPU-only: the C-ABI library loads, exports every symbol include/danbing_b200.h declares, and
refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np

from oracle_abi import ROOT, abi


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "danbing_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(dbt_[a-z0-9_]+)\s*\(", hdr)))


def test_header_and_python_mirror_agree():
    assert declared_symbols() == sorted(abi.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for s in declared_symbols():
        assert hasattr(lib, s), s
    assert lib.dbt_abi_version() == header_facts()["abi"]


_facts = None


def header_facts():
    """sizeof of every struct of include/danbing_b200.h and its version, as a C compiler sees them"""
    global _facts
    if _facts is None:
        import subprocess
        import tempfile
        src = ('#include <stdio.h>\n#include "danbing_b200.h"\nint main(void){printf("abi %d\\nparams %zu\\nindex %zu\\nbatch %zu\\nresult %zu\\n'
               'rawbatch %zu\\nrawinfo %zu\\nrecord %zu\\ncounters %d\\n", DBT_ABI_VERSION, sizeof(dbt_params), sizeof(dbt_index), sizeof(dbt_batch), '
               'sizeof(dbt_batch_result), sizeof(dbt_raw_batch), sizeof(dbt_raw_info), sizeof(dbt_pair_record), DBT_N_COUNTERS);return 0;}\n')
        with tempfile.TemporaryDirectory() as d:
            with open(os.path.join(d, "s.c"), "w") as f:
                f.write(src)
            subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "s"), os.path.join(d, "s.c")], check=True)
            out = subprocess.run([os.path.join(d, "s")], capture_output=True, text=True, check=True).stdout
        _facts = {l.split()[0]: int(l.split()[1]) for l in out.splitlines()}
    return _facts


def test_struct_sizes():
    """the ctypes mirrors (danbing-tk_b200/abi.py) against the header compiled by gcc"""
    f = header_facts()
    assert C.sizeof(abi.Params) == f["params"]
    assert C.sizeof(abi.Index) == f["index"]
    assert C.sizeof(abi.Batch) == f["batch"]
    assert C.sizeof(abi.BatchResult) == f["result"]
    assert C.sizeof(abi.RawBatch) == f["rawbatch"]
    assert C.sizeof(abi.RawInfo) == f["rawinfo"]
    assert abi.DBT_N_COUNTERS == f["counters"]
    assert abi.RECORD_HEADER_DTYPE.itemsize == f["record"] == 64


def test_no_cpu_fallback(lib):
    if lib.dbt_device_count() > 0:
        return
    p = abi.default_params()
    h = C.c_void_p()
    rc = lib.dbt_create(C.byref(h), C.byref(p))
    assert rc == -1 and b"no CPU fallback" in lib.dbt_last_error()


def test_bad_arguments(lib):
    h = C.c_void_p()
    for k, v in (("k", 0), ("k", 28), ("max_read_len", 10), ("max_batch_pairs", 0), ("bucket_slots", 3), ("bucket_slots", 8)):
        p = abi.default_params()
        setattr(p, k, v)
        assert lib.dbt_create(C.byref(h), C.byref(p)) == -3, k
    p = abi.default_params(n_filter=1)
    assert lib.dbt_create(C.byref(h), C.byref(p)) == -3          # -kf 1 x divides by zero in the reference


def test_host_packer_layout(lib):
    """dbt_pack_reads (host side of north_star item 1): 2-bit codes + invalid mask + no-valid-window flag"""
    seqs = ["ACGTACGTACGTACGTACGTACGTA", "ACGTNNNNacgtACGTACGTACGTACGT", "NNNN", "A" * 150]
    bases = np.frombuffer("".join(seqs).encode(), np.uint8).copy()
    off = np.zeros(5, np.uint64)
    off[1:] = np.cumsum([len(s) for s in seqs])
    n = lib.dbt_packed_bytes(150, 2)
    out = np.zeros(n, np.uint8)
    assert lib.dbt_pack_reads(21, 150, 2, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p), 1) == 0
    W = 8
    p = out[: 2 * 2 * W * 8].view("<u8").reshape(2, 2, W)          # [mate][pair][word]
    meta = out[2 * 2 * W * 8:].view("<u2").reshape(2, 2)
    code = {"A": 0, "C": 1, "G": 2, "T": 3}
    for r, s in enumerate(seqs):
        m, pair = r & 1, r >> 1
        rec = p[m, pair]
        for i, ch in enumerate(s):
            inv = (int(rec[5 + i // 64]) >> (i % 64)) & 1
            assert inv == (ch not in code)
            if not inv:
                assert (int(rec[i // 32]) >> (2 * (i % 32))) & 3 == code[ch]
        assert (int(rec[5 + len(s) // 64]) >> (len(s) % 64)) & 1 == 1 or len(s) == 192
        assert int(meta[m, pair]) & 0x7FFF == len(s)
    assert [int(meta[r & 1, r >> 1]) >> 15 for r in range(4)] == [0, 1, 1, 0]


def _numpy_pack(seqs, k, max_len):
    """independent model of the packed record (csrc/kmer.cuh): every word, incl. the zero code under invalid bases"""
    WB, WM = (max_len + 31) // 32, (max_len + 63) // 64
    W = (WB + WM + 1) & ~1
    recs = np.zeros((len(seqs), W), np.uint64)
    meta = np.zeros(len(seqs), np.uint16)
    lut = np.full(256, 4, np.uint8)
    for i, ch in enumerate(b"ACGT"):
        lut[ch] = i
    for r, s in enumerate(seqs):
        c = np.full(WM * 64, 4, np.uint8)
        c[: len(s)] = lut[np.frombuffer(s, np.uint8)]
        bad = c == 4
        codes = np.where(bad, 0, c).astype(np.uint64)[: WB * 32].reshape(WB, 32)
        recs[r, :WB] = (codes << (2 * np.arange(32, dtype=np.uint64))).sum(axis=1, dtype=np.uint64)
        recs[r, WB:WB + WM] = (bad.reshape(WM, 64).astype(np.uint64) << np.arange(64, dtype=np.uint64)).sum(axis=1, dtype=np.uint64)
        run, best = 0, 0
        for b in bad[: len(s)]:
            run = 0 if b else run + 1
            best = max(best, run)
        meta[r] = len(s) | (0 if best >= k else 0x8000)
    return recs, meta


def test_host_packer_random_vs_model(lib, monkeypatch):
    """AVX2 and SWAR packers against the numpy model: all lengths 0..max, N / lower-case / arbitrary bytes, k = 17/21/25"""
    rng = np.random.default_rng(7)
    for max_len, k in ((150, 21), (250, 25), (96, 17), (64, 21)):
        seqs = []
        for L in list(range(0, max_len + 1)) + [max_len] * 50:
            s = rng.choice(np.frombuffer(b"ACGT", np.uint8), L)
            mode = rng.integers(0, 4)
            if mode == 1 and L:
                s[rng.integers(0, L, max(1, L // 30))] = ord("N")
            elif mode == 2 and L:
                s[rng.integers(0, L, max(1, L // 10))] = rng.integers(0, 256, max(1, L // 10))
            elif mode == 3 and L:
                s[rng.integers(0, L)] = ord("acgt"[rng.integers(0, 4)])
            seqs.append(s.astype(np.uint8).tobytes())
        if len(seqs) & 1:
            seqs.append(b"")
        npairs = len(seqs) // 2
        want, wmeta = _numpy_pack(seqs, k, max_len)
        bases = np.frombuffer(b"".join(seqs) + b"\0" * 64, np.uint8).copy()[: sum(map(len, seqs))]
        off = np.zeros(len(seqs) + 1, np.uint64)
        off[1:] = np.cumsum([len(s) for s in seqs])
        W = want.shape[1]
        for swar in (False, True):
            if swar:
                monkeypatch.setenv("DBT_HOST_PACK_SWAR", "1")
            else:
                monkeypatch.delenv("DBT_HOST_PACK_SWAR", raising=False)
            out = np.full(lib.dbt_packed_bytes(max_len, npairs), 0xAB, np.uint8)
            assert lib.dbt_pack_reads(k, max_len, npairs, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(out, abi.u8p), 3) == 0
            p = out[: 2 * npairs * W * 8].view("<u8").reshape(2, npairs, W)
            meta = out[2 * npairs * W * 8:].view("<u2").reshape(2, npairs)
            for m in (0, 1):
                assert np.array_equal(p[m], want[m::2]), (max_len, k, swar, m)
                assert np.array_equal(meta[m], wmeta[m::2]), (max_len, k, swar, m)


def test_device_packer_body_matches_host_packer(tmp_path):
    """csrc/pack_read.cuh (the per-lane SWAR body of pack_kernel, compiled for the host) == dbt_pack_reads' packer:
    every length 0..max, every alignment inside a staged tile, N / lower-case / arbitrary bytes, k = 17/21/25"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "pack_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(root, "tests/native/pack_check.cpp"),
                    os.path.join(root, "danbing-tk_b200/csrc/host_pack.cpp")], check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.startswith("OK"), out.stdout
    ncase, nnovalid, nfast = map(int, out.stdout.split()[1:4])
    assert ncase >= 50000 and 1000 < nnovalid < ncase - 1000
    assert 5000 < nfast < ncase - 5000            # reads the fast (clean-read) path packed itself; it refused exactly the others


def test_header_constants_match_the_python_mirror():
    """every DBT_F_ / DBT_ST_ / DBT_MF_ / DBT_LAYOUT_ / DBT_RAW_ constant of include/danbing_b200.h against danbing-tk_b200/abi.py"""
    hdr = open(os.path.join(ROOT, "include", "danbing_b200.h")).read()
    defs = {m.group(1): int(m.group(2), 0) for m in re.finditer(r"#define\s+DBT_((?:F|ST|MF|LAYOUT|RAW)_[A-Z0-9_]+)\s+(0x[0-9a-fA-F]+|\d+)u?\b", hdr)}
    assert len(defs) >= 30
    checked = 0
    for name, val in defs.items():
        if hasattr(abi, name):
            assert getattr(abi, name) == val, name
            checked += 1
    assert checked >= 25                                  # the mirror carries (nearly) all of them
    for must in ("F_KAM", "F_BAIT", "F_EXTRACT", "ST_COUNTED", "ST_THREAD_FAILED", "MF_BF", "LAYOUT_PACKED", "RAW_SLOWPATH"):
        assert must in defs and hasattr(abi, must), must
