"""numpy readers/writers for the reference's RPGG files and read helpers -- tests/bench plumbing.

File layouts: reference writers src/kmertools.cpp:274-279 (.kmers.dbi),
src/binaryKmerIO.hpp:128-139 (.kdb), :179-188 (.ar); readers src/aQueryFasta_thread.h:654-698.
The product's own loaders are C++ (danbing-tk_b200/host/loaders.cpp); these are independent
restatements used to feed the oracle and to cross-check the C++ loaders.
"""
import gzip
import os

import numpy as np


def _open(fn, mode="rb"):
    if os.path.exists(fn):
        return open(fn, mode)
    if os.path.exists(fn + ".gz"):
        return gzip.open(fn + ".gz", mode)
    raise FileNotFoundError(fn)


def read_dbi(fn):
    with _open(fn) as f:
        b = f.read()
    nk = int(np.frombuffer(b, "<u8", 1, 0)[0])
    keys = np.frombuffer(b, "<u8", nk, 8).copy()
    vals = np.frombuffer(b, "<u4", nk, 8 + 8 * nk).copy()
    o = 8 + 12 * nk
    nvv = int(np.frombuffer(b, "<u8", 1, o)[0])
    vv = np.frombuffer(b, "<u4", nvv, o + 8).copy()
    assert len(b) == o + 8 + 4 * nvv, "kmers.dbi size mismatch"
    return keys, vals, vv


def write_dbi(fn, keys, vals, vv):
    with open(fn, "wb") as f:
        f.write(np.uint64(len(keys)).tobytes())
        f.write(np.ascontiguousarray(keys, "<u8").tobytes())
        f.write(np.ascontiguousarray(vals, "<u4").tobytes())
        f.write(np.uint64(len(vv)).tobytes())
        f.write(np.ascontiguousarray(vv, "<u4").tobytes())


def read_kdb(fn):
    with _open(fn) as f:
        b = f.read()
    nloci = int(np.frombuffer(b, "<u8", 1, 0)[0])
    n = np.frombuffer(b, "<u8", nloci, 8).copy()
    nk = int(np.frombuffer(b, "<u8", 1, 8 + 8 * nloci)[0])
    ks = np.frombuffer(b, "<u8", nk, 16 + 8 * nloci).copy()
    assert int(n.sum()) == nk
    return n, ks


def write_kdb(fn, n, ks):
    with open(fn, "wb") as f:
        f.write(np.uint64(len(n)).tobytes())
        f.write(np.ascontiguousarray(n, "<u8").tobytes())
        f.write(np.uint64(len(ks)).tobytes())
        f.write(np.ascontiguousarray(ks, "<u8").tobytes())


def read_ar(fn):
    with _open(fn) as f:
        b = f.read()
    nk = int(np.frombuffer(b, "<u8", 1, 0)[0])
    return np.frombuffer(b, "<u8", nk, 8).copy()


def read_kmers_text(fn):
    """'>locus' / 'kmer[\\tval]' text -> (n_per_locus u64[nloci], kmers u64[], vals u64[]) in FILE order."""
    n, ks, vs = [], [], []
    with _open(fn, "rt") as f:
        for line in f:
            if line[0] == ">":
                n.append(0)
            else:
                p = line.split()
                ks.append(int(p[0]))
                vs.append(int(p[1]) if len(p) > 1 else 0)
                n[-1] += 1
    return np.array(n, "<u8"), np.array(ks, "<u8"), np.array(vs, "<u8")


def write_kmers_text(fn, n, ks, vs=None):
    with open(fn, "w") as f:
        j = 0
        for l, c in enumerate(n):
            f.write(f">{l}\n")
            for _ in range(int(c)):
                f.write(f"{int(ks[j])}\t{int(vs[j]) if vs is not None else 0}\n")
                j += 1


def load_index(pref, tr_order=None, qc=None):
    """Host view of an RPGG for dbt_load_index / dbo_index_new.

    tr_order: (tr_n, tr_kmers) already in the reference's OUTPUT order (libstdc++ unordered_map
    iteration order). If None, PREF.tr.order.kmers(.gz) (a copy of a reference `-on` output) is
    used when present, else file order of PREF.tr.kmers (fine for everything but the byte order of
    .trkmc.ar)."""
    keys, vals, vv = read_dbi(pref + ".kmers.dbi")
    fl_n, fl_k = read_kdb(pref + ".fl.kdb")
    if tr_order is None:
        for cand in (pref + ".tr.order.kmers", pref + ".tr.kmers"):
            if os.path.exists(cand) or os.path.exists(cand + ".gz"):
                tr_n, tr_k, _ = read_kmers_text(cand)
                break
    else:
        tr_n, tr_k = tr_order
    ix = dict(nloci=len(tr_n), dbi_keys=keys, dbi_vals=vals, vv=vv, tr_n=tr_n, tr_kmers=tr_k, fl_n=fl_n, fl_kmers=fl_k)
    if qc is not None:
        ix["qc"] = np.ascontiguousarray(qc, "u1")
    return ix


# --------------------------------------------------------------------------------------------
# reads
# --------------------------------------------------------------------------------------------
def prune_pe(title):
    """prunePEinfo, src/aQueryFasta_thread.cpp:455-462"""
    if len(title) >= 2 and title[-2] == "/" and title[-1] in "12":
        return title[:-2]
    return title


def pair_fasta(fn, min_read_size=0):
    """Restates the FASTA pairing of CountWords (cpp:1950-1975): returns (titles, seqs) with
    seqs[2i] = late mate ('read 1'), seqs[2i+1] = early mate; pairs with a short mate dropped."""
    titles, seqs, db = [], [], {}
    with _open(fn, "rt") as f:
        while True:
            t = f.readline()
            if not t:
                break
            s = f.readline().rstrip("\n")
            t = prune_pe(t.rstrip("\n"))
            if t in db:
                s2 = db.pop(t)
                if len(s) < min_read_size or len(s2) < min_read_size:
                    continue
                titles.append(t)
                seqs.append(s)
                seqs.append(s2)
            else:
                db[t] = s
    return titles, seqs


def seqs_to_batch(seqs):
    """list of str/bytes -> (bases u8[], offsets u64[n+1])"""
    bs = [s.encode() if isinstance(s, str) else s for s in seqs]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs])
    bases = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(1, np.uint8)
    return bases, off


def write_pairs_tsv(fn, seqs):
    with open(fn, "w") as f:
        for i in range(0, len(seqs), 2):
            f.write(f"{seqs[i]}\t{seqs[i+1]}\n")


# --------------------------------------------------------------------------------------------
# records -> text (mirrors of the reference's writers, for comparing against its stdout)
# --------------------------------------------------------------------------------------------
_CH = "*.="


def unpack_as(as_bytes, nk):
    a = np.asarray(as_bytes, dtype=np.uint8)
    i = np.arange(nk)
    return (a[i >> 2] >> (2 * (i & 3))) & 3


def rle(states):
    """km_asgn_t::annot2str_, cpp:121-138; returns '*' for an empty annotation (cpp:1652)."""
    if len(states) == 0:
        return "*"
    out, ct, a0 = [], 1, int(states[0])
    for a1 in states[1:]:
        a1 = int(a1)
        if a1 != a0:
            out.append(f"{ct}{_CH[a0]}")
            ct = 1
        else:
            ct += 1
        a0 = a1
    out.append(f"{ct}{_CH[a0]}")
    return "".join(out)


def _flag(m, bit):
    return 1 if (int(m["flags"]) & bit) else 0


def trace_line(rec, nloci):
    """One record -> the TSV line oracle/ref_harness.cpp prints (for harness-vs-oracle-vs-GPU diffs)."""
    h = rec["h"]
    st = int(h["status"])
    dst = int(h["dst"]) if st >= 3 else -1
    dst0 = int(h["dst0"]) if st >= 3 else -1
    parts = [str(int(h["pair_index"])), str(st), str(int(h["nhash0"])), str(int(h["nhash1"])), str(dst), str(dst0),
             str(int(h["fc"])), str(int(h["rc"]))]
    for m in range(2):
        mt = h["mate"][m]
        parts.append(":".join(str(x) for x in (_flag(mt, 1), _flag(mt, 2), _flag(mt, 16), _flag(mt, 32), int(mt["si"]),
                                               int(mt["ei"]), int(mt["si_"]), int(mt["ei_"]), int(mt["nt"]),
                                               int(mt["bs"]), int(mt["ti"]))))
        parts.append(rle(unpack_as(rec["as"][m], int(mt["nk"]))))
    return "\t".join(parts)


def kam_line(rec, title, seq1, seq2, qual1=".", qual2=".", src="."):
    """writeKmerAssignments, cpp:1646-1681. title without '>' is printed; mate 2 fields first."""
    h = rec["h"]
    r1, r2 = h["mate"][0], h["mate"][1]
    dst, dst0 = int(h["dst"]), int(h["dst0"])
    dst0s = dst0 if dst0 != dst else -1                      # km_asgn_t::assign, cpp:118

    def na(v):
        return "." if int(v) == -1 else str(int(v))

    def fields(r):
        return ":".join([str(_flag(r, 1)), str(_flag(r, 2)), str(_flag(r, 4)), str(_flag(r, 8)), str(_flag(r, 16)),
                         str(_flag(r, 32)), "0", "0", na(r["si"]), na(r["nt"]), na(r["bs"]), na(r["ti"])])
    as1 = rle(unpack_as(rec["as"][0], int(r1["nk"])))
    as2 = rle(unpack_as(rec["as"][1], int(r2["nk"])))
    return "\t".join([src, str(dst), str(dst0s), str(int(r2["ei"]) - int(r2["si"])), str(int(r1["ei"]) - int(r1["si"])),
                      "kf:hf:bf:qf:af:rm:qn:qm:si:nt:bs:ti", fields(r2), fields(r1), as2, as1, title[1:], seq2, qual2,
                      seq1, qual1])
