"""GPU parity tests (run with `-m gpu` on the B200 box): the CUDA path, called through the C ABI,
against the golden vectors of the unmodified reference and against the CPU oracle on seeded inputs.
Integer / byte work: every comparison is bit-exact."""
import ctypes as C

import numpy as np
import pytest

from common import golden_case, golden_outputs, golden_reads, params_for, summary_text, synthetic_case
from oracle_abi import Oracle, abi, rpgg_io

pytestmark = pytest.mark.gpu

CASES = [("s1", "cth45"), ("s1", "cth10"), ("s1", "kf0"), ("multi", "cth45"), ("multi", "cth10"), ("multi", "kf0")]


def run_gpu(ix, params, seqs, batch=None):
    ctx = abi.Context(params, abi.make_index_struct(ix))
    bases, off = rpgg_io.seqs_to_batch(seqs)
    npairs = len(seqs) // 2
    batch = batch or npairs
    recs_all, cnt = [], np.zeros(abi.DBT_N_COUNTERS, np.uint64)
    for b0 in range(0, max(npairs, 1), batch):
        b1 = min(npairs, b0 + batch)
        o = off[2 * b0: 2 * b1 + 1]
        recs, c = ctx.run(bases, o, first_pair_index=b0, slot=(b0 // batch) % 2)
        recs_all.append(recs.copy())
        cnt += c
    trkmc, kmc, nmap, tot = ctx.download()
    assert np.array_equal(tot, cnt)
    recs = np.concatenate(recs_all) if recs_all else None
    ctx.close()
    return recs, cnt, trkmc, kmc, nmap


@pytest.mark.parametrize("case_name,run_name", CASES)
def test_gpu_matches_reference_golden(case_name, run_name):
    """kam records, .trkmc.ar, .tr.summary.txt and the stderr counters of the reference binary"""
    case = golden_case(case_name)
    run = case["meta"]["runs"][run_name]
    titles, seqs = golden_reads(case, run_name)
    exp = golden_outputs(case, run_name)
    recs, cnt, trkmc, kmc, nmap = run_gpu(case["ix"], params_for(run), seqs)
    assert np.array_equal(trkmc, exp["trkmc"])
    assert summary_text(nmap, kmc) == exp["summary"]
    kam = [rpgg_io.kam_line(r, titles[int(r["h"]["pair_index"])], seqs[2 * int(r["h"]["pair_index"])],
                            seqs[2 * int(r["h"]["pair_index"]) + 1]) for r in recs]
    assert kam == exp["kam"]
    c = dict(zip(abi.COUNTER_NAMES, cnt.tolist()))
    ref, hc = run["counters"], run["harness_counters"]
    assert c["nReads"] == ref["reads processed in total"]
    assert c["nSubFiltered"] == ref["removed by subsampled kmer-filter"]
    assert c["nKmerFiltered"] == ref["removed by kmer-filter"]
    assert c["nLocusAssignFiltered"] == ref["removed during locus assignment"]
    assert c["nThreadingReads"] == ref["entered threading step"] == c["nFeasibleReads"]
    assert c["nAsgnReads"] == ref["assigned to TR region"]
    assert (c["nShort"], c["nhash0"], c["nhash1"]) == (hc["nShort"], hc["nhash0"], hc["nhash1"])
    assert c["ninc"] == int(exp["trkmc"].sum())


@pytest.mark.parametrize("case_name,run_name", [("s1", "cth45"), ("multi", "cth45"), ("multi", "cth10"), ("multi", "kf0")])
def test_gpu_trace_matches_reference_functions(case_name, run_name):
    """every pair's internals (status, probe counts, dst/dst0, top.fc/top.rc at the early stop, flags,
    segment, annotation) vs the reference's own functions driven pair by pair (harness trace)"""
    case = golden_case(case_name)
    run = case["meta"]["runs"][run_name]
    titles, seqs = golden_reads(case, run_name)
    exp = golden_outputs(case, run_name)
    recs, cnt, trkmc, kmc, nmap = run_gpu(case["ix"], params_for(run, flags=abi.F_KAM | abi.F_TRACE), seqs)
    got = [rpgg_io.trace_line(r, case["ix"]["nloci"]) for r in recs]
    assert len(got) == len(exp["trace"])
    bad = [i for i, (a, b) in enumerate(zip(got, exp["trace"])) if a != b]
    assert not bad, f"{len(bad)} pairs differ, first: {bad[0]}\n gpu {got[bad[0]]}\n ref {exp['trace'][bad[0]]}"
    assert np.array_equal(trkmc, exp["trkmc"])


def oracle_vs_gpu(ix, seqs, p_gpu, p_oracle=None, batch=None):
    bases, off = rpgg_io.seqs_to_batch(seqs)
    orecs, ocnt, otr, okmc, onmap = Oracle(ix).run(p_oracle or p_gpu, bases, off)
    recs, cnt, trkmc, kmc, nmap = run_gpu(ix, p_gpu, seqs, batch=batch)
    assert np.array_equal(trkmc, otr)
    assert np.array_equal(kmc, okmc) and np.array_equal(nmap, onmap)
    trace = (p_gpu.flags & abi.F_TRACE) != 0
    if trace:
        assert recs.tobytes() == orecs.tobytes()                       # record buffers byte-identical
    else:
        keep = orecs[orecs["h"]["status"] == 6].copy()
        keep["h"]["nhash0"] = 0                                         # per-pair nhash0, top.fc/top.rc are trace-mode fields
        keep["h"]["fc"] = 0
        keep["h"]["rc"] = 0
        assert recs.tobytes() == keep.tobytes()
    names = [n for n in abi.COUNTER_NAMES if n not in abi.MEASUREMENT_COUNTERS]
    assert {n: int(cnt[i]) for i, n in enumerate(abi.COUNTER_NAMES) if n in names} == \
           {n: int(ocnt[i]) for i, n in enumerate(abi.COUNTER_NAMES) if n in names}
    return recs, cnt


@pytest.mark.parametrize("k,cth,flank,bucket", [(21, 45, 300, 2), (21, 10, 300, 4), (17, 30, 250, 2), (25, 20, 300, 4), (21, 45, 300, 4)])
def test_gpu_matches_oracle_seeded(k, cth, flank, bucket):
    """fresh synthetic RPGG with shared segments + chimeric/borderline/N/short reads; k = 17/21/25"""
    ix, seqs, _ = synthetic_case(nloci=60, seed=200 + k + cth, k=k, flank=flank, npairs=8000, read_seed=cth,
                                 frac_locus=0.4, frac_chimeric=0.3, frac_border=0.15)
    p = abi.default_params(k=k, cth=cth, flags=abi.F_KAM | abi.F_TRACE, max_batch_pairs=8192)
    p.bucket_slots = bucket
    oracle_vs_gpu(ix, seqs, p)


@pytest.mark.parametrize("kf", [(2, 1), (3, 2), (4, 2), (4, 4), (6, 2), (8, 3)], ids=lambda x: f"kf{x[0]}-{x[1]}")
@pytest.mark.parametrize("trace", [True, False], ids=["trace", "production"])
def test_gpu_subfilter_kf_variants(kf, trace):
    """`-kf N M` other than the default 4 1: subfilter_kernel's mate-2 phase by four lanes per pair (N <= 4) and by the per-thread
    loop (N > 4), the break at the M-th hit and the per-pair nhash0 it leaves (compared under DBT_F_TRACE), with and without the
    presence bitmap's shortcut for background k-mers (reads here are a mix of locus, chimeric, borderline and background pairs)"""
    ix, seqs, _ = synthetic_case(nloci=40, seed=500 + 10 * kf[0] + kf[1], k=21, flank=300, npairs=6000, read_seed=kf[0] + kf[1],
                                 frac_locus=0.3, frac_chimeric=0.2, frac_border=0.2, frac_onemate=0.15, err=0.004)
    p = abi.default_params(k=21, cth=30, n_filter=kf[0], nm_filter=kf[1], flags=abi.F_KAM | (abi.F_TRACE if trace else 0), max_batch_pairs=8192)
    recs, cnt = oracle_vs_gpu(ix, seqs, p)
    c = dict(zip(abi.COUNTER_NAMES, cnt.tolist()))
    assert c["nSubFiltered"] > 1000 and c["nAsgnReads"] > 100


@pytest.mark.parametrize("coop", [False, True], ids=["per-lane-staging", "coalesced-staging"])
@pytest.mark.parametrize("k,cth,bucket,err", [(21, 45, 2, 0.0), (21, 30, 4, 0.004), (17, 25, 2, 0.002), (25, 40, 2, 0.0), (21, 10, 2, 0.01)])
def test_gpu_fast_path_matches_oracle(k, cth, bucket, err, coop, monkeypatch):
    """production mode (no DBT_F_TRACE): candidate_fast_kernel decides one-locus pairs in closed form and defers the
    rest to the generic kernel; one-mate pairs sweep the order-dependent band 2*cth <= hits < 3*cth-1.  Both forms of the bucket
    staging: per lane (tables in L2) and one coalesced request per bucket (tables far beyond L2, forced here by DBT_COOP_STAGE)"""
    monkeypatch.setenv("DBT_COOP_STAGE", "1" if coop else "0")
    monkeypatch.setenv("DBT_PREFILTER", "0" if coop else "1")      # genome-scale tables also run without subfilter's presence bitmap
    ix, seqs, _ = synthetic_case(nloci=60, seed=300 + k + cth, k=k, flank=300, npairs=12000, read_seed=cth + 1,
                                 frac_locus=0.35, frac_chimeric=0.15, frac_border=0.1, frac_onemate=0.3, err=err)
    p = abi.default_params(k=k, cth=cth, flags=abi.F_KAM, max_batch_pairs=12288)
    p.bucket_slots = bucket
    recs, cnt = oracle_vs_gpu(ix, seqs, p)
    c = dict(zip(abi.COUNTER_NAMES, cnt.tolist()))
    assert c["nLocusAssignFiltered"] > 0 and c["nKmerFiltered"] > 0 and c["nAsgnReads"] > 0


def test_gpu_fast_path_long_reads():
    """250 bp reads (MAXJ = 8 instantiation) through the fast path"""
    ix, seqs, _ = synthetic_case(nloci=30, seed=77, k=21, flank=400, npairs=4000, read_seed=3, rlen=250, flen=600,
                                 frac_locus=0.5, frac_chimeric=0.1, frac_border=0.1, frac_onemate=0.2)
    p = abi.default_params(k=21, cth=45, flags=abi.F_KAM, max_read_len=250, max_batch_pairs=4096)
    oracle_vs_gpu(ix, seqs, p)


def test_gpu_batches_and_slots_accumulate():
    """same reads in 5 batches over both slots == one batch (counts are order-independent sums)"""
    ix, seqs, _ = synthetic_case(nloci=40, seed=5, npairs=5000, read_seed=9, frac_locus=0.5, frac_chimeric=0.2, frac_border=0.1)
    p = abi.default_params(k=21, cth=30, flags=abi.F_KAM, max_batch_pairs=1024)
    oracle_vs_gpu(ix, seqs, p, batch=1000)


def test_gpu_ka_and_qc_flags():
    """-ka (no kam records; assignTRkmc skips removed mates, cpp:1451) and -qc (cpp:2059-2062)"""
    ix, seqs, _ = synthetic_case(nloci=40, seed=6, npairs=4000, read_seed=10, frac_locus=0.5, frac_chimeric=0.2, frac_border=0.1)
    qc = (np.arange(ix["nloci"]) % 3 != 0).astype(np.uint8)
    ix = dict(ix, qc=qc)
    p = abi.default_params(k=21, cth=30, flags=abi.F_QC | abi.F_TRACE, max_batch_pairs=4096)
    recs, cnt = oracle_vs_gpu(ix, seqs, p)
    assert int(cnt[abi.COUNTER_NAMES.index("nQCFiltered")]) > 0
    p2 = abi.default_params(k=21, cth=30, flags=0, max_batch_pairs=4096)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    ctx = abi.Context(p2, abi.make_index_struct(ix))
    recs, c = ctx.run(bases, off)
    assert len(recs) == 0                                              # -ka: nothing on stdout
    o = Oracle(ix).run(p2, bases, off)
    assert np.array_equal(ctx.download()[0], o[2])


@pytest.mark.parametrize("flags", [abi.F_KAM | abi.F_TRACE | abi.F_THREADING | abi.F_CORRECTION, abi.F_KAM | abi.F_THREADING])
def test_gpu_threading_flag_head_semantics(flags):
    """-g/-gc at the reference's HEAD: pairs enter the threading step and none passes (cpp:2070-2090)"""
    ix, seqs, _ = synthetic_case(nloci=30, seed=8, npairs=3000, read_seed=12, frac_locus=0.5, frac_chimeric=0.2, frac_onemate=0.1)
    p = abi.default_params(k=21, cth=30, flags=flags, max_batch_pairs=4096)
    recs, cnt = oracle_vs_gpu(ix, seqs, p)
    c = dict(zip(abi.COUNTER_NAMES, cnt.tolist()))
    assert c["nThreadingReads"] > 0 and c["nFeasibleReads"] == 0 and c["nAsgnReads"] == 0 and c["ninc"] == 0


def test_gpu_empty_and_degenerate_batches():
    ix, seqs, _ = synthetic_case(nloci=10, seed=7, npairs=50, read_seed=11)
    p = abi.default_params(k=21, cth=10, flags=abi.F_KAM | abi.F_TRACE, max_batch_pairs=64)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    recs, cnt = ctx.run(np.zeros(1, np.uint8), np.zeros(1, np.uint64))     # empty batch
    assert len(recs) == 0 and int(cnt.sum()) == 0
    weird = ["N" * 150, "ACGT" * 30, "A" * 20, "ACGTACGTAC" * 15, "acgt" * 37, "ACGT" * 37, "", "ACGT" * 37]
    bases, off = rpgg_io.seqs_to_batch(weird)
    recs, cnt = ctx.run(bases, off)
    orecs = Oracle(ix).run(p, bases, off)[0]
    assert recs.tobytes() == orecs.tobytes()
    too_long = ["A" * 151, "C" * 150]
    bases, off = rpgg_io.seqs_to_batch(too_long)
    with pytest.raises(abi.DbtError, match="capacity"):
        ctx.run(bases, off)
    ctx.close()


def test_gpu_packed_layout_equals_ascii(lib):
    """host packer + DBT_LAYOUT_PACKED upload == device pack kernel"""
    ix, seqs, _ = synthetic_case(nloci=30, seed=8, npairs=3000, read_seed=12, frac_locus=0.5, frac_chimeric=0.2)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    npairs = len(seqs) // 2
    p = abi.default_params(k=21, cth=30, flags=abi.F_KAM | abi.F_TRACE, max_batch_pairs=4096)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    r1, c1 = ctx.run(bases, off)
    r1 = r1.copy()
    t1 = ctx.download()[0].copy()
    ctx.reset()
    packed = np.zeros(lib.dbt_packed_bytes(150, npairs), np.uint8)
    assert lib.dbt_pack_reads(21, 150, npairs, abi.ptr(bases, abi.u8p), abi.ptr(off, abi.u64p), abi.ptr(packed, abi.u8p), 4) == 0
    b = ctx.make_batch(packed, None, layout=abi.LAYOUT_PACKED, npairs=npairs)
    rec = np.zeros(npairs * ctx.stride, np.uint8)
    r = abi.BatchResult()
    r.capacity = npairs
    r.records = abi.ptr(rec, abi.u8p)
    abi.check(lib, lib.dbt_submit(ctx.h, C.byref(b), C.byref(r), 0), "submit")
    abi.check(lib, lib.dbt_wait(ctx.h, 0), "wait")
    assert rec.tobytes() == r1.tobytes()
    assert np.array_equal(ctx.download()[0], t1)
    ctx.close()


def test_gpu_index_consistency_is_checked():
    ix, seqs, _ = synthetic_case(nloci=10, seed=9, npairs=10, read_seed=13)
    bad = dict(ix)
    bad["tr_kmers"] = ix["tr_kmers"].copy()
    bad["tr_kmers"][0] ^= np.uint64(0x155)                # a TR k-mer that kmers.dbi does not hold
    p = abi.default_params(k=21, cth=10, max_batch_pairs=64)
    with pytest.raises(abi.DbtError, match="index"):
        abi.Context(p, abi.make_index_struct(bad))


def _vv_max_list(ix):
    vv, m, o = ix["vv"], 0, 0
    while o < len(vv):
        m = max(m, int(vv[o]))
        o += int(vv[o]) + 1
    return m


@pytest.mark.parametrize("spill_at", [None, 2])
def test_gpu_vote_spill_pool(spill_at, monkeypatch):
    """k-mers listed in > 100 loci (one segment dropped into 130 loci): a pair's locus vote touches more loci than the
    shared-memory table holds and spills into the locked global pool (candidate.cuh VoteHits).  Generic kernel for every
    pair (TRACE) and fast + generic kernels (production), then BOTH slots in flight at once sharing the pool;
    DBT_VOTE_SPILL_AT=2 makes nearly every multi-locus vote spill."""
    if spill_at is not None:
        monkeypatch.setenv("DBT_VOTE_SPILL_AT", str(spill_at))
    ix, seqs, _ = synthetic_case(nloci=160, seed=91, npairs=6000, read_seed=17, multi_share=(130, 140),
                                 frac_locus=0.5, frac_chimeric=0.25, frac_border=0.1)
    assert _vv_max_list(ix) >= 100
    for flags in (abi.F_KAM | abi.F_TRACE, abi.F_KAM):
        p = abi.default_params(k=21, cth=30, flags=flags, max_batch_pairs=8192)
        oracle_vs_gpu(ix, seqs, p)
    # two batches in flight on the two slots (the CLI's worker loop): counts equal the oracle's on the concatenation
    p = abi.default_params(k=21, cth=30, flags=abi.F_KAM | abi.F_TRACE, max_batch_pairs=4096)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    _, _, otr, okmc, onmap = Oracle(ix).run(p, bases, off)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    L = ctx.L
    npairs, half = len(seqs) // 2, len(seqs) // 4
    keep = []
    for rep in range(3):
        ctx.reset()
        for s, (b0, b1) in enumerate(((0, half), (half, npairs))):
            rec = np.zeros((b1 - b0) * ctx.stride, np.uint8)
            r = abi.BatchResult()
            r.capacity = b1 - b0
            r.records = abi.ptr(rec, abi.u8p)
            b = ctx.make_batch(bases, off[2 * b0: 2 * b1 + 1], first_pair_index=b0)
            keep.append((rec, r, b))
            abi.check(L, L.dbt_submit(ctx.h, C.byref(b), C.byref(r), s), "submit")
        for s in (0, 1):
            abi.check(L, L.dbt_wait(ctx.h, s), "wait")
        trkmc, kmc, nmap, _ = ctx.download()
        assert np.array_equal(trkmc, otr) and np.array_equal(kmc, okmc) and np.array_equal(nmap, onmap)
    ctx.close()
