#!/usr/bin/env python3
"""Generates tests/golden/* by running the UNMODIFIED reference (oracle/_ref, built from
/root/reference/src by oracle/Makefile) -- run in the build container only:

    make -C oracle ref && python tests/golden/make_golden.py

Cases
  s1     : SURVEY.md §8(c) recipe on the reference's own test/input haplotype FASTAs (config 1):
           fa2kmers -k 21 -fsi 700 -fso 700, ktools serialize, sim_reads tiling reads (sub-sampled)
           plus trap reads (N, lower case, short mates, corrupted tails, background)
  multi  : 40-locus synthetic RPGG with shared segments (multi-locus k-mers -> the vv path) built
           with the reference's fa2kmers + ktools, reads incl. chimeric two-locus pairs (trap 5)
For each case and flag set the reference binary's outputs are stored: OUT.trkmc.ar, OUT.tr.summary.txt,
stdout kam lines, the final stderr counters, the `-on` k-mer order, and the per-pair trace of
oracle/_ref/ref_harness (the reference's own functions called pair by pair).
"""
import gzip
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_abi import REF_DIR, pkg, rpgg_io, run_ref_binary, run_ref_harness  # noqa: E402

synth = pkg.synth
REF = "/root/reference"
RUNS = [dict(name="cth45", k=21, cth=45, kf=(4, 1), c=40), dict(name="cth10", k=21, cth=10, kf=(4, 1), c=40),
        dict(name="kf0", k=21, cth=30, kf=(0, 0), c=20)]


def gz_write(fn, data):
    with gzip.GzipFile(fn, "wb", mtime=0) as f:
        f.write(data if isinstance(data, bytes) else data.encode())


def counters_from_stderr(err):
    keys = ["reads processed in total", "removed by subsampled kmer-filter", "removed by kmer-filter", "removed by bait locus",
            "removed by qual filter", "removed during locus assignment", "removed by QC filter", "entered threading step",
            "passsed threading", "assigned to TR region"]
    out = {}
    for line in err.splitlines():
        for k in keys:
            if k in line:
                out[k] = int(line.split()[0])
    return out


def finish_case(case_dir, tmp, pref, reads, k_runs):
    """reads: list of (title, mate_first, mate_second); pref: RPGG prefix in tmp."""
    fa = os.path.join(tmp, "reads.fa")
    synth.write_fasta(fa, reads)
    os.makedirs(case_dir, exist_ok=True)
    for ext in (".kmers.dbi", ".fl.kdb", ".tre.kdb"):
        gz_write(os.path.join(case_dir, "pan" + ext + ".gz"), open(pref + ext, "rb").read())
    gz_write(os.path.join(case_dir, "pan.tr.kmers.gz"), open(pref + ".tr.kmers", "rb").read())
    gz_write(os.path.join(case_dir, "reads.fa.gz"), open(fa, "rb").read())
    meta = {"libstdcxx": subprocess.run("echo '#include <bits/c++config.h>\n__GLIBCXX__' | g++ -E -x c++ - | tail -1", shell=True,
                                        capture_output=True, text=True).stdout.strip(), "runs": {}}
    for run in k_runs:
        out = os.path.join(tmp, "out_" + run["name"])
        kam, err = run_ref_binary(pref, fa, out, k=run["k"], cth=run["cth"], kf=run["kf"], c=run["c"])
        run_ref_binary(pref, fa, out + "n", k=run["k"], cth=run["cth"], kf=run["kf"], c=run["c"], named=True)
        titles, seqs = rpgg_io.pair_fasta(fa, run["cth"] + run["k"] - 1)
        trace, htrk, hsum, hc = run_ref_harness(pref, seqs, os.path.join(tmp, "h_" + run["name"]), run["k"], run["cth"],
                                                run["kf"][0], run["kf"][1], run["c"])
        trk = open(out + ".trkmc.ar", "rb").read()
        assert rpgg_io.read_ar(out + ".trkmc.ar").tolist() == htrk.tolist(), "harness glue disagrees with the reference binary"
        assert open(out + ".tr.summary.txt").read() == hsum
        d = os.path.join(case_dir, run["name"])
        os.makedirs(d, exist_ok=True)
        gz_write(os.path.join(d, "out.trkmc.ar.gz"), trk)
        open(os.path.join(d, "out.tr.summary.txt"), "w").write(open(out + ".tr.summary.txt").read())
        gz_write(os.path.join(d, "out.kam.gz"), kam)
        gz_write(os.path.join(d, "harness.trace.tsv.gz"), "\n".join(trace) + "\n")
        meta["runs"][run["name"]] = dict(run, counters=counters_from_stderr(err), harness_counters=hc,
                                         sha256_trkmc=hashlib.sha256(trk).hexdigest(), sha256_kam=hashlib.sha256(kam.encode()).hexdigest(),
                                         npairs=len(titles))
    # `-on` output order = libstdc++ unordered_map iteration order (SURVEY.md F4 trap 6)
    tn, tk, _ = rpgg_io.read_kmers_text(os.path.join(tmp, "out_" + k_runs[0]["name"] + "n.tr.kmers"))
    gz_write(os.path.join(case_dir, "pan.tr.order.u64.gz"), np.concatenate([np.array([len(tn)], "<u8"), tn, tk]).tobytes())
    json.dump(meta, open(os.path.join(case_dir, "meta.json"), "w"), indent=1, sort_keys=True)


def case_s1(tmp):
    I = os.path.join(REF, "test/input")
    haps = [os.path.join(I, h + ".fa") for h in ("HG00514.0", "HG00514.1", "HG00733.0", "HG00733.1")]
    pref = os.path.join(tmp, "pan")
    run = lambda *a: subprocess.run(list(a), check=True, capture_output=True)
    run(REF_DIR + "/fa2kmers", "-k", "21", "-fsi", "700", "-fso", "700", "-on", pref, "-fa", "4", *haps)
    run(REF_DIR + "/fa2kmers", "-tr", "-k", "22", "-fsi", "700", "-fso", "700", "-on", pref + "22", "-fa", "4", *haps)
    os.rename(pref + "22.tr.kmers", pref + ".tre.kmers")
    run(REF_DIR + "/ktools", "serialize", pref)
    seqs = []
    for h in haps:
        s = "".join(l.strip() for l in open(h).read().splitlines()[1:]).upper()
        seqs.append(s.encode())
    loci = [seqs]            # one locus, four haplotypes
    rng = np.random.default_rng(11)
    reads = []
    for hi, s in enumerate(seqs):      # sim_reads-style tiling, step 40
        for beg in range(0, len(s) - 500 + 1, 40):
            reads.append((">hap%d:%d-%d" % (hi, beg, beg + 500), s[beg:beg + 150], synth.revcomp(s[beg + 350:beg + 500])))
    reads += synth.sim_reads(loci, 1200, seed=5, frac_locus=0.3, frac_chimeric=0.0, frac_border=0.4, frac_n=0.1, frac_short=0.03,
                             frac_lower=0.03)
    order = rng.permutation(len(reads))
    reads = [reads[i] for i in order]
    return pref, reads


def case_multi(tmp):
    loci = synth.make_loci(40, seed=7, nhap=4, flank=300, share_frac=0.3, share_len=120, tr_median=200)
    fns = synth.write_hap_fastas(loci, os.path.join(tmp, "hap"))
    pref = os.path.join(tmp, "pan")
    run = lambda *a: subprocess.run(list(a), check=True, capture_output=True)
    run(REF_DIR + "/fa2kmers", "-k", "21", "-fsi", "300", "-fso", "300", "-on", pref, "-fa", str(len(fns)), *fns)
    run(REF_DIR + "/fa2kmers", "-tr", "-k", "22", "-fsi", "300", "-fso", "300", "-on", pref + "22", "-fa", str(len(fns)), *fns)
    os.rename(pref + "22.tr.kmers", pref + ".tre.kmers")
    run(REF_DIR + "/ktools", "serialize", pref)
    reads = synth.sim_reads(loci, 4000, seed=3, frac_locus=0.35, frac_chimeric=0.35, frac_border=0.15)
    return pref, reads


if __name__ == "__main__":
    for name, fn in (("s1", case_s1), ("multi", case_multi)):
        with tempfile.TemporaryDirectory() as tmp:
            pref, reads = fn(tmp)
            d = os.path.join(HERE, name)
            shutil.rmtree(d, ignore_errors=True)
            finish_case(d, tmp, pref, reads, RUNS)
            print(name, "done:", sum(os.path.getsize(os.path.join(r, f)) for r, _, fs in os.walk(d) for f in fs), "bytes")
