"""Shared helpers for the graph-threading (P1) parity tests: seeded RPGG + graph + reads with
substitution / indel / N errors, the reference harness in threading mode, the C restatement per read."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

from oracle_abi import Oracle, abi, oracle_lib, pkg, ref_bin, rpgg_io

synth = pkg.synth


class ThreadOut(C.Structure):
    _fields_ = [("rc", C.c_int), ("err", C.c_int), ("nkm", C.c_uint32), ("kmers", C.c_uint64 * 1024),
                ("nes", C.c_uint32), ("es_t", C.c_uint8 * 1024), ("es_g", C.c_uint8 * 1024),
                ("ntr", C.c_uint32), ("tr", C.c_uint8 * 1024)]


def thread_case(nloci=8, seed=7, nhap=3, flank=200, k=21, npairs=400, rseed=1, sub=0.01, ins=0.002, dele=0.002,
                frac_n=0.05, frac_bg=0.0, rlen=150, outdir=None, write_text=True):
    """-> (pref, ix incl. graph, seqs in internal mate order, loci)."""
    d = outdir or tempfile.mkdtemp(prefix="dbt_thr_")
    os.makedirs(d, exist_ok=True)
    loci = synth.make_loci(nloci, seed=seed, nhap=nhap, flank=flank, share_frac=0.3, share_len=80, tr_median=200)
    pref = os.path.join(d, "pan")
    ix = synth.build_rpgg(loci, pref, k=k, fsi=flank, fso=flank, write_text=write_text)
    ix.update(synth.build_graph(loci, k, pref if write_text else None))
    rng = np.random.default_rng(rseed)
    seqs = []
    for _ in range(npairs):
        if rng.random() < frac_bg:
            a, b = synth.random_read(rng, rlen), synth.random_read(rng, rlen)
        else:
            l = int(rng.integers(0, nloci))
            hap = loci[l][int(rng.integers(0, nhap))]
            a, b = synth.sim_pair(hap, rng, rlen, min(400, len(hap)))
            a = synth.mutate_indel(a, rng, sub, ins, dele)
            b = synth.mutate_indel(b, rng, sub, ins, dele)
            if rng.random() < frac_n:
                aa = bytearray(a)
                aa[int(rng.integers(0, rlen))] = ord("N")
                a = bytes(aa)
        seqs.append(b.decode())
        seqs.append(a.decode())
    return pref, ix, seqs, loci


def thread_params(k=21, cth=30, thread=50, corr=1, maxnc=4, aln=0, extra_flags=0, **kw):
    flags = abi.F_THREADING | (abi.F_CORRECTION if corr else 0) | extra_flags
    flags |= {0: 0, 1: abi.F_ALN, 2: abi.F_ALN | abi.F_ALN_MIN}[aln]
    p = abi.default_params(k=k, cth=cth, flags=flags, **kw)
    p.thread_cth, p.max_ncorrection = thread, maxnc
    return p


def run_ref_thread_harness(pref, seqs, outpref, p, aln=1):
    """Reference harness in threading mode -> (trace rows, aln text, trkmc, counters)."""
    tsv = outpref + ".pairs.tsv"
    rpgg_io.write_pairs_tsv(tsv, seqs)
    corr = 1 if p.flags & abi.F_CORRECTION else 0
    cmd = [ref_bin("ref_harness"), pref, tsv, outpref, str(p.k), str(p.cth), str(p.n_filter), str(p.nm_filter), str(p.nm_tr), "0",
           f"thread={p.thread_cth}", f"corr={corr}", f"maxnc={p.max_ncorrection}", f"aln={aln}"]
    r = subprocess.run(cmd, check=True, capture_output=True, text=True)
    rows = [l.rstrip("\n").split("\t") for l in open(outpref + ".trace.tsv")]
    trkmc = rpgg_io.read_ar(outpref + ".trkmc.ar")
    counters = dict((l.split("\t")[0], int(l.split("\t")[1])) for l in open(outpref + ".counters.txt").read().splitlines())
    return rows, r.stdout, trkmc, counters


def parse_thread_row(f):
    """trace row of a pair that entered threading -> (idx, status, dst_before, [(alned, es, tr, kmers)]*2)."""
    idx, status = int(f[0]), int(f[1])
    if len(f) < 21:
        return idx, status, None, None
    dst = int(f[12])
    mates = []
    for m in range(2):
        alned, es, tr, km = f[13 + 4 * m: 17 + 4 * m]
        kms = [] if km == "." else [int(x) & 0xFFFFFFFFFFFFFFFF for x in km.split(",")]
        mates.append((int(alned), es, tr, kms))
    return idx, status, dst, mates


def oracle_thread_read(o, locus, seq, p):
    """C restatement (oracle/dbt_oracle_thread.c) -> (rc, err, es string, tr string, kmers list)."""
    L = oracle_lib()
    L.dbo_thread_read.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p, C.c_uint32, C.POINTER(abi.Params), C.POINTER(ThreadOut)]
    out = ThreadOut()
    s = seq.encode()
    rc = L.dbo_thread_read(o.h, locus, s, len(s), C.byref(p), C.byref(out))
    es = "".join(chr(out.es_t[i]) + (chr(out.es_g[i]) if out.es_g[i] else "_") for i in range(out.nes))
    tr = bytes(out.tr[:out.ntr]).decode()
    return rc, out.err, es, tr, [int(out.kmers[i]) for i in range(out.nkm)]


def es_string(es_bytes, n):
    return "".join(abi.ES_T[b & 7] + abi.ES_G[(b >> 3) & 7] for b in es_bytes[:n])


def tr_string(tr_bytes, n):
    return "".join(abi.TR_C[b] for b in tr_bytes[:n])


def check_records_vs_harness(recs, rows, nloci):
    """Record-level parity of dbt_aln records (GPU or oracle, one per pair, input order) with the harness trace."""
    n = 0
    for f in rows:
        idx, status, dst, mates = parse_thread_row(f)
        r = recs[idx]
        if mates is None:
            assert int(r["h"]["status"]) == status, f"pair {idx}"
            continue
        want_status = abi.ST_COUNTED if status == 6 else abi.ST_THREAD_FAILED
        assert int(r["h"]["status"]) == want_status, f"pair {idx}"
        assert int(r["h"]["dst"]) == (dst if status == 6 else nloci), f"pair {idx}"
        for m in range(2):
            alned, es, tr, kms = mates[m]
            a = r["aln"][m]
            assert int(a["h"]["err"]) == 0
            assert int(a["h"]["alned"]) == alned, f"pair {idx} mate {m}"
            assert es_string(a["es"], int(a["h"]["nes"])) == es, f"pair {idx} mate {m}"
            assert tr_string(a["tr"], int(a["h"]["ntr"])) == tr, f"pair {idx} mate {m}"
            assert int(a["h"]["nkm"]) == len(kms) and int(a["h"]["kmers_hash"]) == abi.fnv1a_u64(kms), f"pair {idx} mate {m}"
        n += 1
    return n


def ref_tr_order(pref, tmpdir, k=21):
    """(tr_n, tr_kmers) in the reference's output order = libstdc++ unordered_map iteration order
    (SURVEY.md F4 trap 6), read off a `-on` run of the unmodified reference binary on one dummy pair."""
    from oracle_abi import run_ref_binary
    fa = os.path.join(tmpdir, "dummy.fa")
    with open(fa, "w") as f:
        f.write(">d/1\n" + "A" * 60 + "\n>d/2\n" + "C" * 60 + "\n")
    out = os.path.join(tmpdir, "order")
    run_ref_binary(pref, fa, out, k=k, named=True)
    tn, tk, _ = rpgg_io.read_kmers_text(out + ".tr.kmers")
    return tn, tk
