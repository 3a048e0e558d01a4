"""Shared helpers for the parity tests."""
import gzip
import json
import os

import numpy as np

from oracle_abi import Oracle, abi, pkg, rpgg_io  # noqa: F401

synth = pkg.synth
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_case(name):
    """-> dict(ix=host index in reference output order, titles, seqs (reference pairing incl. short mates), meta, dir)"""
    d = os.path.join(GOLDEN, name)
    pref = os.path.join(d, "pan")
    o = np.frombuffer(gzip.open(pref + ".tr.order.u64.gz").read(), "<u8")
    nloci = int(o[0])
    tr_n, tr_k = o[1:1 + nloci].copy(), o[1 + nloci:].copy()
    ix = rpgg_io.load_index(pref, tr_order=(tr_n, tr_k))
    meta = json.load(open(os.path.join(d, "meta.json")))
    return dict(ix=ix, meta=meta, dir=d, fasta=os.path.join(d, "reads.fa"))


def golden_reads(case, run):
    r = case["meta"]["runs"][run]
    titles, seqs = rpgg_io.pair_fasta(case["fasta"], r["cth"] + r["k"] - 1)
    return titles, seqs


def golden_outputs(case, run):
    d = os.path.join(case["dir"], run)
    return dict(trkmc=rpgg_io.read_ar(os.path.join(d, "out.trkmc.ar")),
                summary=open(os.path.join(d, "out.tr.summary.txt")).read(),
                kam=gzip.open(os.path.join(d, "out.kam.gz"), "rt").read().splitlines(),
                trace=gzip.open(os.path.join(d, "harness.trace.tsv.gz"), "rt").read().splitlines())


def params_for(run, flags=abi.F_KAM, max_read_len=150, max_batch_pairs=1 << 16, **kw):
    p = abi.default_params(k=run["k"], cth=run["cth"], n_filter=run["kf"][0], nm_filter=run["kf"][1], nm_tr=run["c"],
                           flags=flags, max_read_len=max_read_len, max_batch_pairs=max_batch_pairs)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def summary_text(nmap, kmc):
    return "".join(f"{int(a)}\t{int(b)}\n" for a, b in zip(nmap, kmc))


def synthetic_case(nloci=60, seed=21, k=21, nhap=4, flank=300, npairs=6000, read_seed=5, multi_share=None, **simkw):
    """RPGG + mixed trap reads generated here (numpy), for oracle-vs-GPU parity beyond the goldens."""
    import tempfile
    loci = synth.make_loci(nloci, seed=seed, nhap=nhap, flank=flank, share_frac=0.3, share_len=120, tr_median=200, multi_share=multi_share)
    tmp = tempfile.mkdtemp(prefix="dbt_syn_")
    ix = synth.build_rpgg(loci, os.path.join(tmp, "syn"), k=k, fsi=flank, fso=flank, write_text=False)
    reads = synth.sim_reads(loci, npairs, seed=read_seed, **simkw)
    seqs = []
    for t, a, b in reads:            # arrival order /1 then /2 -> internal order (late mate first)
        seqs.append(b.decode())
        seqs.append(a.decode())
    return ix, seqs, loci
