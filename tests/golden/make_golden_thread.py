"""Generates tests/golden/thread: outputs of the REFERENCE's isThreadFeasible (through oracle/_ref/ref_harness,
threading mode, decision D1 glue) on a seeded RPGG + graph + reads with sub/indel/N errors.
Run in the build container (needs /root/reference-built oracle/_ref): python tests/golden/make_golden_thread.py"""
import gzip
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_abi import abi  # noqa: E402
import thread_common as tc  # noqa: E402

CASE = dict(seed=31, rseed=9, sub=0.03, ins=0.008, dele=0.008, npairs=600, thread=40, maxnc=4, cth=25, frac_n=0.1, frac_bg=0.1)


def main():
    out = os.path.join(HERE, "thread")
    os.makedirs(out, exist_ok=True)
    tmp = tempfile.mkdtemp()
    gk = {k: v for k, v in CASE.items() if k not in ("thread", "corr", "maxnc", "cth")}
    pk = {k: CASE[k] for k in ("thread", "corr", "maxnc", "cth", "k") if k in CASE}
    pref, ix, seqs, _ = tc.thread_case(outdir=tmp, **gk)
    p = tc.thread_params(**pk)
    rows, aln, trkmc, counters = tc.run_ref_thread_harness(pref, seqs, os.path.join(tmp, "ref"), p, aln=1)
    _, aln_min, _, _ = tc.run_ref_thread_harness(pref, seqs, os.path.join(tmp, "ref2"), p, aln=2)
    hist, n = {}, 0
    for f in rows:
        _, _, _, mates = tc.parse_thread_row(f)
        if mates:
            for m in mates:
                hist[str(m[0])] = hist.get(str(m[0]), 0) + 1
                n += 1
    with open(os.path.join(tmp, "ref.trace.tsv"), "rb") as f, gzip.GzipFile(os.path.join(out, "harness.trace.tsv.gz"), "wb", mtime=0) as g:
        shutil.copyfileobj(f, g)
    with gzip.GzipFile(os.path.join(out, "out.aln.gz"), "wb", mtime=0) as g:
        g.write(aln.encode())
    with gzip.GzipFile(os.path.join(out, "out.aln_min.gz"), "wb", mtime=0) as g:
        g.write(aln_min.encode())
    with gzip.GzipFile(os.path.join(out, "out.trkmc.ar.gz"), "wb", mtime=0) as g:
        g.write(np.uint64(len(trkmc)).tobytes() + trkmc.astype("<u8").tobytes())
    meta = dict(case=CASE, counters=counters, threaded_reads=n, rc_hist=hist,
                reads_fnv=abi.fnv1a_u64(np.frombuffer("".join(seqs).encode(), np.uint8).astype("<u8")),
                note="trkmc is in PREF.tr.kmers FILE order of the synthetic RPGG as loaded by the reference harness "
                     "(unordered_map iteration order); tests compare per-locus multisets or use tr order from the harness run")
    json.dump(meta, open(os.path.join(out, "meta.json"), "w"), indent=1)
    print(meta)


if __name__ == "__main__":
    main()
