"""Pins the C restatement of graph threading (oracle/dbt_oracle_thread.c) against the reference's own
isThreadFeasible (src/aQueryFasta_thread.cpp:1114-1260) driven read by read by oracle/_ref/ref_harness,
and against the committed golden trace (tests/golden/thread, made by tests/golden/make_golden_thread.py)."""
import gzip
import json
import os

import numpy as np
import pytest

from oracle_abi import Oracle, abi, have_ref, rpgg_io
import thread_common as tc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "thread")

CASES = [
    dict(seed=11, rseed=2, sub=0.03, ins=0.01, dele=0.01, npairs=500, thread=30, maxnc=4),
    dict(seed=12, rseed=3, sub=0.05, ins=0.02, dele=0.02, npairs=500, thread=20, maxnc=8, cth=20),
    dict(seed=13, rseed=4, sub=0.02, ins=0.0, dele=0.0, npairs=400, thread=85, maxnc=3),
    dict(seed=14, rseed=5, sub=0.0, ins=0.02, dele=0.02, npairs=500, thread=40, maxnc=6, frac_n=0.3),
    dict(seed=15, rseed=6, sub=0.02, ins=0.005, dele=0.005, npairs=400, thread=50, corr=0),
    dict(seed=16, rseed=7, sub=0.04, ins=0.01, dele=0.01, npairs=500, thread=30, maxnc=4, k=17, cth=20),
    dict(seed=17, rseed=8, sub=0.04, ins=0.01, dele=0.01, npairs=500, thread=30, maxnc=4, k=25, cth=20),
]


def split_case(c):
    pk = {k: c[k] for k in ("thread", "corr", "maxnc", "cth", "k") if k in c}
    gk = {k: v for k, v in c.items() if k not in ("thread", "corr", "maxnc", "cth")}
    return gk, pk


def check_rows(rows, seqs, ix, p):
    o = Oracle(ix)
    nread, hist = 0, {}
    for f in rows:
        idx, status, dst, mates = tc.parse_thread_row(f)
        if mates is None:
            continue
        for m in range(2):
            alned, es, tr, kms = mates[m]
            rc, err, mes, mtr, mkm = tc.oracle_thread_read(o, dst, seqs[2 * idx + m], p)
            assert not err
            assert (rc, mes, mtr, mkm) == (alned, es, tr, kms), f"pair {idx} mate {m}"
            nread += 1
            hist[rc] = hist.get(rc, 0) + 1
    return nread, hist


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("ci", range(len(CASES)))
def test_restatement_vs_reference_isThreadFeasible(ci, tmp_path):
    gk, pk = split_case(CASES[ci])
    pref, ix, seqs, _ = tc.thread_case(outdir=str(tmp_path), **gk)
    p = tc.thread_params(**pk)
    rows, _, _, _ = tc.run_ref_thread_harness(pref, seqs, str(tmp_path / "ref"), p)
    nread, hist = check_rows(rows, seqs, ix, p)
    assert nread > 100
    if pk.get("corr", 1):
        assert hist.get(2, 0) > 0          # reads threaded WITH edits were exercised


def test_restatement_vs_golden_trace():
    meta = json.load(open(os.path.join(GOLD, "meta.json")))
    gk, pk = split_case(meta["case"])
    _, ix, seqs, _ = tc.thread_case(write_text=False, **gk)
    assert abi.fnv1a_u64(np.frombuffer("".join(seqs).encode(), np.uint8).astype("<u8")) == meta["reads_fnv"]
    p = tc.thread_params(**pk)
    rows = [l.rstrip("\n").split("\t") for l in gzip.open(os.path.join(GOLD, "harness.trace.tsv.gz"), "rt")]
    nread, hist = check_rows(rows, seqs, ix, p)
    assert nread == meta["threaded_reads"] and {str(k): v for k, v in hist.items()} == meta["rc_hist"]


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("ci", [0, 3, 4])
def test_pair_level_D1_glue_vs_harness(ci, tmp_path):
    """dbo_run in restored-threading mode: statuses, records, counters and per-locus counts equal the harness run."""
    gk, pk = split_case(CASES[ci])
    pref, ix, seqs, _ = tc.thread_case(outdir=str(tmp_path), frac_bg=0.2, **gk)
    p = tc.thread_params(aln=1, **pk)
    rows, _, ref_trkmc, ref_cnt = tc.run_ref_thread_harness(pref, seqs, str(tmp_path / "ref"), p)
    # the harness dumps counts in the reference's unordered_map order: load the index in that order
    ix2 = rpgg_io.load_index(pref, tr_order=tc.ref_tr_order(pref, str(tmp_path)))
    ix2.update({k: ix[k] for k in ("g_n", "g_kmers", "g_masks")})
    bases, off = rpgg_io.seqs_to_batch(seqs)
    recs, cnt, trkmc, kmc, nmap = Oracle(ix2).run(p, bases, off)
    assert tc.check_records_vs_harness(recs, rows, int(ix["nloci"])) > 50
    assert np.array_equal(trkmc, ref_trkmc)
    names = dict(zip(abi.COUNTER_NAMES, [int(x) for x in cnt]))
    for key in ("nReads", "nShort", "nSubFiltered", "nKmerFiltered", "nLocusAssignFiltered", "nThreadingReads", "nFeasibleReads", "nhash0", "nhash1"):
        assert names[key] == ref_cnt[key], key
    assert kmc.sum() == 0 and nmap.sum() == 0
