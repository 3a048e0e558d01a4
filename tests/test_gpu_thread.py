"""GPU parity of restored graph threading (P1: thread_kernel through the C ABI) against the oracle restatement,
the reference's own isThreadFeasible (oracle/_ref/ref_harness) and the committed golden trace."""
import gzip
import json
import os

import numpy as np
import pytest

from oracle_abi import Oracle, abi, have_ref, rpgg_io
import thread_common as tc
from test_oracle_thread import CASES, GOLD, split_case

pytestmark = pytest.mark.gpu


def gpu_run(ix, p, seqs, batch=None):
    bases, off = rpgg_io.seqs_to_batch(seqs)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    assert ctx.threading
    recs, cnt = ctx.run(bases, off)
    trkmc, kmc, nmap, tot = ctx.download()
    ctx.close()
    return recs, cnt, trkmc, kmc, nmap


def assert_same_as_oracle(ix, p, seqs):
    bases, off = rpgg_io.seqs_to_batch(seqs)
    orecs, ocnt, otr, okmc, onmap = Oracle(ix).run(p, bases, off)
    recs, cnt, trkmc, kmc, nmap = gpu_run(ix, p, seqs)
    assert len(recs) == len(orecs)
    bad = [i for i in range(len(recs)) if recs[i].tobytes() != orecs[i].tobytes()]
    if bad:
        i = bad[0]
        g, o = recs[i], orecs[i]
        msg = [f"{len(bad)} records differ; first pair {i}: status gpu {g['h']['status']} oracle {o['h']['status']} dst {g['h']['dst']}/{o['h']['dst']}"]
        for m in range(2):
            ga, oa = g["aln"][m], o["aln"][m]
            msg.append(f" mate {m}: alned {ga['h']['alned']}/{oa['h']['alned']} err {ga['h']['err']} nes {ga['h']['nes']}/{oa['h']['nes']} ntr {ga['h']['ntr']}/{oa['h']['ntr']} "
                       f"nkm {ga['h']['nkm']}/{oa['h']['nkm']} hash_eq {ga['h']['kmers_hash'] == oa['h']['kmers_hash']}")
            msg.append("  gpu es " + tc.es_string(ga["es"], int(ga["h"]["nes"])))
            msg.append("  ora es " + tc.es_string(oa["es"], int(oa["h"]["nes"])))
            msg.append("  gpu tr " + tc.tr_string(ga["tr"], int(ga["h"]["ntr"])))
            msg.append("  ora tr " + tc.tr_string(oa["tr"], int(oa["h"]["ntr"])))
        msg.append(" seqs " + seqs[2 * i] + " " + seqs[2 * i + 1])
        raise AssertionError("\n".join(msg))
    assert np.array_equal(trkmc, otr) and np.array_equal(kmc, okmc) and np.array_equal(nmap, onmap)
    for i, name in enumerate(abi.COUNTER_NAMES):
        if name not in abi.MEASUREMENT_COUNTERS:
            assert int(cnt[i]) == int(ocnt[i]), name
    return recs, cnt


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_trace_records_counts_counters_vs_oracle(ci):
    gk, pk = split_case(CASES[ci])
    _, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.2, **gk)
    p = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=4096, **pk)
    recs, cnt = assert_same_as_oracle(ix, p, seqs)
    st = recs["h"]["status"]
    assert (st == abi.ST_COUNTED).sum() > 20


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("ci", [0, 1, 3])
def test_trace_records_vs_reference_harness(ci, tmp_path):
    gk, pk = split_case(CASES[ci])
    pref, ix, seqs, _ = tc.thread_case(outdir=str(tmp_path), frac_bg=0.2, **gk)
    p = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=4096, **pk)
    rows, _, _, ref_cnt = tc.run_ref_thread_harness(pref, seqs, str(tmp_path / "ref"), p)
    recs, cnt, trkmc, _, _ = gpu_run(ix, p, seqs)
    assert tc.check_records_vs_harness(recs, rows, int(ix["nloci"])) > 20
    names = dict(zip(abi.COUNTER_NAMES, [int(x) for x in cnt]))
    for key in ("nThreadingReads", "nFeasibleReads", "nhash0", "nhash1", "nSubFiltered", "nKmerFiltered", "nLocusAssignFiltered"):
        assert names[key] == ref_cnt[key], key


def test_trace_records_vs_golden_trace():
    meta = json.load(open(os.path.join(GOLD, "meta.json")))
    gk, pk = split_case(meta["case"])
    _, ix, seqs, _ = tc.thread_case(write_text=False, **gk)
    p = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=4096, **pk)
    rows = [l.rstrip("\n").split("\t") for l in gzip.open(os.path.join(GOLD, "harness.trace.tsv.gz"), "rt")]
    recs, cnt, trkmc, _, _ = gpu_run(ix, p, seqs)
    assert tc.check_records_vs_harness(recs, rows, int(ix["nloci"])) * 2 == meta["threaded_reads"]
    names = dict(zip(abi.COUNTER_NAMES, [int(x) for x in cnt]))
    for key in ("nThreadingReads", "nFeasibleReads", "nhash0", "nhash1"):
        assert names[key] == meta["counters"][key], key
    gold = np.frombuffer(gzip.open(os.path.join(GOLD, "out.trkmc.ar.gz")).read(), "<u8")[1:]
    assert int(trkmc.sum()) == int(gold.sum())


@pytest.mark.parametrize("aln", [0, 1, 2])
def test_record_selection_follows_a_and_ae(aln):
    """-a: every pair that entered threading; -ae: threaded pairs only; neither: no records (cpp:2232-2239)."""
    gk, pk = split_case(CASES[1])
    _, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.2, **gk)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    pt = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=4096, **pk)
    orecs, ocnt, otr, _, _ = Oracle(ix).run(pt, bases, off)
    st = orecs["h"]["status"]
    want = {0: np.zeros(len(st), bool), 1: (st == abi.ST_COUNTED) | (st == abi.ST_THREAD_FAILED), 2: st == abi.ST_COUNTED}[aln]
    p = tc.thread_params(aln=aln, max_batch_pairs=4096, **pk)
    recs, cnt, trkmc, _, _ = gpu_run(ix, p, seqs)
    assert np.array_equal(trkmc, otr)
    assert [int(x) for x in recs["h"]["pair_index"]] == [int(i) for i in np.nonzero(want)[0]]
    for r in recs:
        o = orecs[int(r["h"]["pair_index"])]
        assert r["h"]["status"] == o["h"]["status"] and r["h"]["dst"] == o["h"]["dst"]
        for m in range(2):
            a, b = r["aln"][m], o["aln"][m]
            assert a["h"]["alned"] == b["h"]["alned"] and a["es"].tobytes() == b["es"].tobytes() and a["tr"].tobytes() == b["tr"].tobytes()


def test_threading_flag_without_graph_keeps_head_behaviour():
    """No graph in the index: `-gc` behaves like the reference's HEAD (nothing passes, cpp:2072-2090)."""
    gk, pk = split_case(CASES[0])
    _, ix, seqs, _ = tc.thread_case(write_text=False, **gk)
    ix = {k: v for k, v in ix.items() if not k.startswith("g_")}
    p = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=4096, **pk)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    assert not ctx.threading
    recs, cnt = ctx.run(bases, off)
    trkmc, _, _, _ = ctx.download()
    assert trkmc.sum() == 0 and (recs["h"]["status"] == abi.ST_NOT_THREADED).sum() > 50


@pytest.mark.parametrize("cfg", [
    dict(seed=41, rseed=1, sub=0.02, ins=0.004, dele=0.004, npairs=6000, thread=60, maxnc=4, nloci=30),
    dict(seed=42, rseed=2, sub=0.06, ins=0.02, dele=0.02, npairs=6000, thread=15, maxnc=12, cth=15, nloci=12),
    dict(seed=43, rseed=3, sub=0.01, ins=0.03, dele=0.03, npairs=6000, thread=25, maxnc=6, cth=20, frac_n=0.2, nloci=12, k=17),
    dict(seed=44, rseed=4, sub=0.03, ins=0.01, dele=0.01, npairs=4000, thread=30, maxnc=5, cth=20, rlen=250, nloci=12),
])
def test_stress_vs_oracle(cfg):
    """many reads, heavy error mixes, 250 bp reads (MAXJ = 8 instantiation): record-exact against the oracle"""
    gk, pk = split_case(cfg)
    _, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.1, **gk)
    p = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=8192, max_read_len=gk.get("rlen", 150), **pk)
    recs, cnt = assert_same_as_oracle(ix, p, seqs)
    assert (recs["h"]["status"] == abi.ST_COUNTED).sum() > 200


@pytest.mark.parametrize("aln", [1, 2])
def test_pinned_result_buffer_gets_the_same_records(aln):
    """dbt_wait copies the device-compacted records straight into a pinned caller buffer (dbt_host_alloc) and stages
    them through its own pinned buffer otherwise; both must return the same records in pair order, over two slots,
    for `-a` (every pair that entered threading) and `-ae` (threaded pairs only)."""
    import ctypes as C
    _, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.3, npairs=600, rseed=11, sub=0.02)
    p = tc.thread_params(aln=aln, max_batch_pairs=1024)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    L = ctx.L
    want, _ = ctx.run(bases, off)                                   # pageable numpy buffer: staged path
    want = want.copy()
    pt = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=1024)
    st = Oracle(ix).run(pt, bases, off)[0]["h"]["status"]
    sel = (st == abi.ST_COUNTED) | (st == abi.ST_THREAD_FAILED) if aln == 1 else st == abi.ST_COUNTED
    assert [int(x) for x in want["h"]["pair_index"]] == [int(i) for i in np.nonzero(sel)[0]]
    assert 0 < len(want) < len(seqs) // 2
    ctx.reset()
    cap = len(seqs) // 2
    for slot in (0, 1):
        buf, owner = abi.pinned_array(L, cap * ctx.stride)
        buf[:] = 0xEE
        r = abi.BatchResult()
        r.capacity = cap
        r.records = abi.ptr(buf, abi.u8p)
        b = ctx.make_batch(bases, off)
        abi.check(L, L.dbt_submit(ctx.h, C.byref(b), C.byref(r), slot), "submit")
        abi.check(L, L.dbt_wait(ctx.h, slot), "wait")
        n = int(r.n_records)
        assert n == len(want)
        assert buf[: n * ctx.stride].tobytes() == want.tobytes()
        assert (buf[n * ctx.stride:] == 0xEE).all()                  # nothing written past the emitted records
        del buf, owner
    ctx.close()


def test_result_capacity_overflow_is_reported():
    _, ix, seqs, _ = tc.thread_case(write_text=False, npairs=300, rseed=12)
    p = tc.thread_params(aln=1, max_batch_pairs=512)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    with pytest.raises(abi.DbtError, match="result buffer holds 5 records"):
        ctx.run(bases, off, capacity=5)
    ctx.close()


def test_mixed_packed_and_ascii_sub_batches_over_both_slots(lib=None):
    """bench.py's hybrid input path: sub-batches alternate between host-packed (dbt_pack_reads + DBT_LAYOUT_PACKED) and ASCII
    uploads on the two slots; records, counts and counters equal one ASCII submission (threading mode, reads with N / errors)."""
    import ctypes as C
    L = abi.load_library()
    _, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.2, nloci=8, seed=21, npairs=1500, frac_n=0.2)
    seqs[10] = seqs[10].lower()
    p = tc.thread_params(aln=1, max_batch_pairs=4096)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    npairs = len(seqs) // 2
    ctx = abi.Context(p, abi.make_index_struct(ix))
    r_all, c_all = ctx.run(bases, off)
    r_all = r_all.copy()
    t_all = [a.copy() for a in ctx.download()[:3]]
    ctx.reset()
    sb = 400
    nsub = (npairs + sb - 1) // sb
    res = [abi.BatchResult() for _ in range(2)]
    bufs = [np.zeros(sb * ctx.stride, np.uint8) for _ in range(2)]
    keep, got, pending = [], [], [None, None]
    for j in range(nsub):
        s = j & 1
        if pending[s] is not None:
            abi.check(L, L.dbt_wait(ctx.h, s), "wait")
            got.append(bufs[s][: int(res[s].n_records) * ctx.stride].copy())
        b0, b1 = j * sb, min(npairs, (j + 1) * sb)
        o = off[2 * b0: 2 * b1 + 1]
        if j % 2:
            packed = np.zeros(L.dbt_packed_bytes(p.max_read_len, b1 - b0), np.uint8)
            assert L.dbt_pack_reads(p.k, p.max_read_len, b1 - b0, abi.ptr(bases, abi.u8p), abi.ptr(o, abi.u64p), abi.ptr(packed, abi.u8p), 2) == 0
            b = ctx.make_batch(packed, None, first_pair_index=b0, layout=abi.LAYOUT_PACKED, npairs=b1 - b0)
            keep.append(packed)
        else:
            b = ctx.make_batch(bases, o, first_pair_index=b0)
        keep.append(b)
        res[s].capacity = sb
        res[s].records = abi.ptr(bufs[s], abi.u8p)
        abi.check(L, L.dbt_submit(ctx.h, C.byref(b), C.byref(res[s]), s), "submit")
        pending[s] = j
    order = sorted((j, s) for s, j in enumerate(pending) if j is not None)
    for j, s in order:
        abi.check(L, L.dbt_wait(ctx.h, s), "wait")
        got.append(bufs[s][: int(res[s].n_records) * ctx.stride].copy())
    assert np.concatenate(got).tobytes() == r_all.tobytes()
    for a, b in zip(ctx.download()[:3], t_all):
        assert np.array_equal(a, b)
    ctx.close()


STRESS = [
    dict(seed=41, rseed=1, sub=0.02, ins=0.004, dele=0.004, npairs=6000, thread=60, maxnc=4, nloci=30),
    dict(seed=42, rseed=2, sub=0.06, ins=0.02, dele=0.02, npairs=6000, thread=15, maxnc=12, cth=15, nloci=12),
    dict(seed=43, rseed=3, sub=0.01, ins=0.03, dele=0.03, npairs=6000, thread=25, maxnc=6, cth=20, frac_n=0.2, nloci=12, k=17),
    dict(seed=44, rseed=4, sub=0.03, ins=0.01, dele=0.01, npairs=4000, thread=30, maxnc=5, cth=20, rlen=250, nloci=12),
    dict(seed=45, rseed=5, sub=0.004, ins=0.0002, dele=0.0002, npairs=8000, thread=85, maxnc=3, nloci=40),     # the bench's error model
]


@pytest.mark.parametrize("merge", [True, False, "coop"], ids=["graph-in-index", "two-tables", "graph-in-index-coalesced-staging"])
@pytest.mark.parametrize("cfg", list(CASES) + STRESS, ids=lambda c: f"s{c.get('seed')}")
def test_production_mode_records_vs_oracle(cfg, merge, monkeypatch):
    """`-a` without DBT_F_TRACE = the production kernels: with the graphs merged into the index candidate_fast_kernel threads the pairs
    itself (thread_fast_kernel only sees what the generic kernel defers), with DBT_NO_GRAPH_MERGE thread_fast_kernel probes the graph
    table; both must give the oracle's records (status, locus, alned, CIGAR, annotation) and counts."""
    if not merge:
        monkeypatch.setenv("DBT_NO_GRAPH_MERGE", "1")
    if merge == "coop":                              # the bucket staging used when the tables are far beyond L2 (stage_buckets_coop)
        monkeypatch.setenv("DBT_COOP_STAGE", "1")
    gk, pk = split_case(cfg)
    _, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.15, **gk)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    rl = gk.get("rlen", 150)
    pt = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=8192, max_read_len=rl, **pk)
    orecs, ocnt, otr, okmc, onmap = Oracle(ix).run(pt, bases, off)
    p = tc.thread_params(aln=1, max_batch_pairs=8192, max_read_len=rl, **pk)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    assert ctx.L.dbt_graph_merged(ctx.h) == (1 if merge else 0)
    recs, cnt = ctx.run(bases, off)
    trkmc, kmc, nmap, _ = ctx.download()
    ctx.close()
    assert np.array_equal(trkmc, otr) and np.array_equal(kmc, okmc) and np.array_equal(nmap, onmap)
    st = orecs["h"]["status"]
    want = np.nonzero((st == abi.ST_COUNTED) | (st == abi.ST_THREAD_FAILED))[0]
    assert [int(x) for x in recs["h"]["pair_index"]] == [int(i) for i in want]
    for r in recs:
        o = orecs[int(r["h"]["pair_index"])]
        assert r["h"]["status"] == o["h"]["status"] and r["h"]["dst"] == o["h"]["dst"], int(r["h"]["pair_index"])
        for m in range(2):
            a, b = r["aln"][m], o["aln"][m]
            assert a["h"]["alned"] == b["h"]["alned"] and a["h"]["nes"] == b["h"]["nes"] and a["h"]["ntr"] == b["h"]["ntr"], (int(r["h"]["pair_index"]), m)
            assert a["es"].tobytes() == b["es"].tobytes() and a["tr"].tobytes() == b["tr"].tobytes(), (int(r["h"]["pair_index"]), m)
    for i, name in enumerate(abi.COUNTER_NAMES):
        if name not in abi.MEASUREMENT_COUNTERS:
            assert int(cnt[i]) == int(ocnt[i]), name
