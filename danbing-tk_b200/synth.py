"""Seeded synthetic RPGGs and reads (SURVEY.md §8(d) S2/S3 shapes) -- tests/bench workload plumbing.

build_rpgg() restates what the reference's `fa2kmers -k K -fsi F -fso F` (src/fa2kmers.cpp:131-167,
buildNuKmers src/aQueryFasta_thread.h:185-213) + `ktools serialize` (src/kmertools.cpp:221-345,
readKmerIndex src/kmerIO.hpp:47-78) produce from haplotype sequences, and writes the reference's file
formats directly; tests/test_synth.py checks it against the real tools.
"""
import os

import numpy as np

from . import rpgg_io

_CODE = np.full(256, 255, np.uint8)
for _i, _c in enumerate(b"ACGT"):
    _CODE[_c] = _i
_ASCII = np.frombuffer(b"ACGT", np.uint8)
_COMP = np.zeros(256, np.uint8)
for _a, _b in zip(b"ACGTN", b"TGCAN"):
    _COMP[_a] = _b


def canonical_kmers(codes, k):
    """codes: u8 array of 0..3 -> canonical k-mers of every window (u64[len-k+1])."""
    n = len(codes) - k + 1
    if n <= 0:
        return np.zeros(0, np.uint64)
    f = np.zeros(n, np.uint64)
    r = np.zeros(n, np.uint64)
    c = codes.astype(np.uint64)
    for j in range(k):
        w = c[j:j + n]
        f |= w << np.uint64(2 * (k - 1 - j))
        r |= (np.uint64(3) - w) << np.uint64(2 * j)
    return np.minimum(f, r)


def make_loci(nloci, seed=20261104, nhap=8, flank=700, share_frac=0.05, share_len=200, tr_median=300,
              tr_min=50, tr_max=5000, sub_rate=0.02, multi_share=None):
    """Returns haps[locus][hap] = ASCII bytes (flank + TR + flank). A fraction of loci copy a segment
    of another locus' TR/flank so that multi-locus k-mers (the vv path) occur.
    multi_share = (n, seg_len): ONE random segment is additionally dropped into n loci, so that its k-mers are
    listed in n loci (vv lists of hundreds of entries, as low-complexity k-mers are in real RPGGs)."""
    loci = []
    for l in range(nloci):
        rng = np.random.default_rng([seed, l])
        lf = rng.integers(0, 4, flank, dtype=np.uint8)
        rf = rng.integers(0, 4, flank, dtype=np.uint8)
        mlen = int(rng.integers(6, 61))
        motif = rng.integers(0, 4, mlen, dtype=np.uint8)
        haps = []
        for h in range(nhap):
            trlen = int(np.clip(rng.lognormal(np.log(tr_median), 0.6), tr_min, tr_max))
            ncopy = max(1, trlen // mlen)
            tr = np.tile(motif, ncopy)
            mut = rng.random(len(tr)) < sub_rate
            tr = np.where(mut, (tr + rng.integers(1, 4, len(tr), dtype=np.uint8)) % 4, tr).astype(np.uint8)
            haps.append(np.concatenate([lf, tr, rf]))
        loci.append(haps)
    rng = np.random.default_rng([seed, 10 ** 9])
    nshare = int(round(nloci * share_frac)) if nloci > 1 else 0
    for _ in range(nshare):
        a, b = rng.choice(nloci, 2, replace=False)
        src = loci[b][0]
        s0 = int(rng.integers(0, max(1, len(src) - share_len)))
        seg = src[s0:s0 + share_len]
        for h in range(nhap):
            d = loci[a][h]
            # drop the shared segment into the TR/flank boundary region of locus a
            p0 = int(min(max(flank - share_len // 2, 0), len(d) - len(seg)))
            d[p0:p0 + len(seg)] = seg
    if multi_share:
        n, seg_len = multi_share
        seg = rng.integers(0, 4, seg_len, dtype=np.uint8)
        for a in rng.choice(nloci, min(n, nloci), replace=False):
            for h in range(nhap):
                d = loci[a][h]
                p0 = int(min(max(flank + 10, 0), len(d) - len(seg)))
                d[p0:p0 + len(seg)] = seg
    return [[_ASCII[h].tobytes() for h in haps] for haps in loci]


def write_hap_fastas(loci, pref):
    """One FASTA per haplotype, one record per locus (fa2kmers input convention). Returns file names."""
    nhap = len(loci[0])
    fns = []
    for h in range(nhap):
        fn = f"{pref}.hap{h}.fa"
        with open(fn, "wb") as f:
            for l, haps in enumerate(loci):
                f.write(b">locus%d.hap%d\n" % (l, h))
                f.write(haps[h] + b"\n")
        fns.append(fn)
    return fns


def locus_kmer_sets(haps, k, fsi, fso):
    """(tr_kmers, fl_kmers) unique canonical k-mer arrays of one locus; fa2kmers.cpp:139-156."""
    tr, fl = [], []
    for s in haps:
        codes = _CODE[np.frombuffer(s, np.uint8)]
        assert codes.max() < 4, "synthetic haplotypes contain only ACGT"
        km = canonical_kmers(codes, k)
        n = len(s)
        tr.append(km[fsi: n - k - fsi + 1])
        if fso:
            fl.append(km[fsi - fso: fsi])
            fl.append(km[n - fsi - (k - 1): n - k - (fsi - fso) + 1])
    tr = np.unique(np.concatenate(tr)) if tr else np.zeros(0, np.uint64)
    fl = np.unique(np.concatenate(fl)) if fl else np.zeros(0, np.uint64)
    return tr, fl


def build_dbi(tr_n, tr_k, fl_n, fl_k):
    """readKmerIndex over .tr.kmers then .fl.kmers (kio:47-78) + vv flattening (kmertools.cpp:243-258).
    List order inside vv = order of first sighting: TR loci ascending, then flank-only loci ascending."""
    tr_l = np.repeat(np.arange(len(tr_n), dtype=np.uint64), tr_n.astype(np.int64))
    fl_l = np.repeat(np.arange(len(fl_n), dtype=np.uint64), fl_n.astype(np.int64))
    km = np.concatenate([tr_k, fl_k])
    lo = np.concatenate([tr_l, fl_l])
    src = np.concatenate([np.zeros(len(tr_k), np.uint8), np.ones(len(fl_k), np.uint8)])
    order = np.lexsort((lo, src, km))           # by k-mer, then TR before FL, then locus
    km, lo, src = km[order], lo[order], src[order]
    # drop (kmer, locus) duplicates (flank sighting of a pair already seen in TR)
    pair_order = np.lexsort((src, lo, km))
    km2, lo2 = km[pair_order], lo[pair_order]
    first = np.ones(len(km2), bool)
    first[1:] = (km2[1:] != km2[:-1]) | (lo2[1:] != lo2[:-1])
    keep = np.zeros(len(km), bool)
    keep[pair_order[first]] = True
    km, lo = km[keep], lo[keep]
    keys, start, cnt = np.unique(km, return_index=True, return_counts=True)
    vals = np.zeros(len(keys), np.uint32)
    single = cnt == 1
    vals[single] = (lo[start[single]] << np.uint64(1)).astype(np.uint32)
    vv = []
    off = 0
    multi = np.nonzero(~single)[0]
    for i in multi:
        n = int(cnt[i])
        vals[i] = (off << 1) | 1
        vv.append(n)
        vv.extend(int(x) for x in lo[start[i]: start[i] + n])
        off += n + 1
    return keys.astype("<u8"), vals.astype("<u4"), np.array(vv, "<u4")


def forward_kmers(codes, k):
    """codes: u8 array of 0..3 -> strand-specific k-mers of every window (u64[len-k+1])."""
    n = len(codes) - k + 1
    if n <= 0:
        return np.zeros(0, np.uint64)
    f = np.zeros(n, np.uint64)
    c = codes.astype(np.uint64)
    for j in range(k):
        f |= c[j:j + n] << np.uint64(2 * (k - 1 - j))
    return f


def locus_graph(haps, k):
    """(nodes u64 sorted, masks u8) of one locus: fa2kmers -g == buildKmerGraph over the WHOLE haplotype
    record, both strands, self-loops dropped (src/fa2kmers.cpp:158, src/aQueryFasta_thread.h:215-243).
    Bit b of the mask: successor ((node & rmask) << 2) + b exists."""
    ns, ms = [], []
    for s in haps:
        codes = _CODE[np.frombuffer(s, np.uint8)]
        for c in (codes, (3 - codes[::-1]).astype(np.uint8)):
            f = forward_kmers(c, k)
            if not len(f):
                continue
            m = np.zeros(len(f), np.uint8)
            nxt = c[k:].astype(np.uint8)
            m[:-1] = np.where(f[:-1] != f[1:], np.uint8(1) << nxt, np.uint8(0))
            ns.append(f)
            ms.append(m)
    if not ns:
        return np.zeros(0, "<u8"), np.zeros(0, np.uint8)
    ns, ms = np.concatenate(ns), np.concatenate(ms)
    order = np.argsort(ns, kind="stable")
    ns, ms = ns[order], ms[order]
    nodes, start = np.unique(ns, return_index=True)
    masks = np.bitwise_or.reduceat(ms, start)
    return nodes.astype("<u8"), masks.astype(np.uint8)


_PAR = {}


def _par_sets(args):
    l0, l1, k, fsi, fso = args
    return [locus_kmer_sets(h, k, fsi, fso) for h in _PAR["loci"][l0:l1]]


def _par_graph(args):
    l0, l1, k = args
    return [locus_graph(h, k) for h in _PAR["loci"][l0:l1]]


def _pmap(fn, loci, extra, nproc):
    """per-locus map over forked workers (the loci are inherited copy-on-write); serial below 4000 loci or with nproc <= 1"""
    if nproc <= 1 or len(loci) < 4000:
        _PAR["loci"] = loci
        return fn((0, len(loci)) + extra)
    import multiprocessing as mp
    _PAR["loci"] = loci
    step = max(64, len(loci) // (nproc * 8))
    jobs = [(a, min(len(loci), a + step)) + extra for a in range(0, len(loci), step)]
    with mp.get_context("fork").Pool(nproc) as pool:
        parts = pool.map(fn, jobs)
    return [x for p in parts for x in p]


_wlib = None


def write_kmers_text_fast(fn, n, ks, vs=None):
    """rpgg_io.write_kmers_text through libdbt_synth.so's C writer when it is built (values: u8 masks or zeros)"""
    global _wlib
    so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libdbt_synth.so")
    if _wlib is None and os.path.exists(so):
        import ctypes as C
        _wlib = C.CDLL(so)
        if not hasattr(_wlib, "dbt_write_kmers_text"):
            _wlib = False
        else:
            _wlib.dbt_write_kmers_text.argtypes = [C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    if not _wlib:
        return rpgg_io.write_kmers_text(fn, n, ks, vs)
    n = np.ascontiguousarray(n, "<u8")
    ks = np.ascontiguousarray(ks, "<u8")
    v = None if vs is None else np.ascontiguousarray(vs, np.uint8)
    rc = _wlib.dbt_write_kmers_text(fn.encode(), len(n), n.ctypes.data, ks.ctypes.data, None if v is None else v.ctypes.data)
    if rc:
        raise OSError(f"dbt_write_kmers_text({fn}) failed: {rc}")


def build_graph(loci, k=21, pref=None, nproc=1):
    """-> dict(g_n, g_kmers, g_masks); with pref also writes PREF.graph.kmers (text, `node<TAB>mask`)."""
    gs = _pmap(_par_graph, loci, (k,), nproc)
    g_n = np.array([len(n) for n, _ in gs], "<u8")
    g_k = np.concatenate([n for n, _ in gs]).astype("<u8") if gs else np.zeros(0, "<u8")
    g_m = np.concatenate([m for _, m in gs]).astype(np.uint8) if gs else np.zeros(0, np.uint8)
    if pref is not None:
        write_kmers_text_fast(pref + ".graph.kmers", g_n, g_k, g_m)
    return dict(g_n=g_n, g_kmers=g_k, g_masks=g_m)


def build_rpgg(loci, pref, k=21, fsi=700, fso=700, write_text=True, nproc=1):
    """Writes PREF.tr.kmers, PREF.fl.kmers (text), PREF.kmers.dbi, PREF.fl.kdb, PREF.tre.kdb (valid, empty).
    Returns the host index dict (tr order = file order)."""
    trs, fls = zip(*_pmap(_par_sets, loci, (k, fsi, fso), nproc))
    tr_n = np.array([len(t) for t in trs], "<u8")
    fl_n = np.array([len(t) for t in fls], "<u8")
    tr_k = np.concatenate(trs).astype("<u8") if len(trs) else np.zeros(0, "<u8")
    fl_k = np.concatenate(fls).astype("<u8") if len(fls) else np.zeros(0, "<u8")
    keys, vals, vv = build_dbi(tr_n, tr_k, fl_n, fl_k)
    os.makedirs(os.path.dirname(os.path.abspath(pref)), exist_ok=True)
    if write_text:
        write_kmers_text_fast(pref + ".tr.kmers", tr_n, tr_k)
        write_kmers_text_fast(pref + ".fl.kmers", fl_n, fl_k)
    rpgg_io.write_dbi(pref + ".kmers.dbi", keys, vals, vv)
    rpgg_io.write_kdb(pref + ".fl.kdb", fl_n, fl_k)
    rpgg_io.write_kdb(pref + ".tre.kdb", np.zeros(len(tr_n), "<u8"), np.zeros(0, "<u8"))
    return dict(nloci=len(tr_n), dbi_keys=keys, dbi_vals=vals, vv=vv, tr_n=tr_n, tr_kmers=tr_k, fl_n=fl_n, fl_kmers=fl_k)


# --------------------------------------------------------------------------------------------
# reads (small-scale numpy simulator for tests; bench.py uses the CUDA generator in csrc/synth.cu)
# --------------------------------------------------------------------------------------------
def revcomp(s):
    return _COMP[np.frombuffer(s, np.uint8)][::-1].tobytes()


def sim_pair(hap, rng, rlen=150, flen=500):
    """error-free pair like src/sim_reads.cpp:74-104: forward read + reverse-complemented far end."""
    beg = int(rng.integers(0, len(hap) - flen + 1))
    return hap[beg:beg + rlen], revcomp(hap[beg + flen - rlen: beg + flen])


def mutate(s, rng, rate):
    a = np.frombuffer(s, np.uint8).copy()
    m = rng.random(len(a)) < rate
    a[m] = _ASCII[rng.integers(0, 4, int(m.sum()))]
    return a.tobytes()


def mutate_indel(s, rng, sub=0.0, ins=0.0, dele=0.0, rlen=None):
    """substitutions (to a DIFFERENT base), insertions and deletions; result cut/padded back to rlen."""
    a = np.frombuffer(s, np.uint8)
    out = bytearray()
    u = rng.random((len(a), 3))
    for i, c in enumerate(a):
        if u[i, 0] < dele:
            continue
        if u[i, 1] < ins:
            out.append(int(_ASCII[rng.integers(0, 4)]))
        if u[i, 2] < sub and c in b"ACGT":
            c = b"ACGT"[(b"ACGT".index(c) + int(rng.integers(1, 4))) % 4]
        out.append(int(c))
    rlen = len(a) if rlen is None else rlen
    while len(out) < rlen:
        out.append(int(_ASCII[rng.integers(0, 4)]))
    return bytes(out[:rlen])


def random_read(rng, rlen=150):
    return _ASCII[rng.integers(0, 4, rlen)].tobytes()


def sim_reads(loci, npairs, seed=1, frac_locus=0.5, frac_chimeric=0.1, frac_border=0.1, frac_n=0.05,
              frac_short=0.01, frac_lower=0.01, err=0.0, rlen=150, frac_onemate=0.0, flen=500):
    """Mixed test workload: list of (title, mate_first, mate_second) in ARRIVAL order (/1 then /2).
    Includes the trap cases of SURVEY.md F4: chimeric two-locus pairs (trap 5), corrupted tails,
    N / lower-case bases (trap 1-2), short mates (trap 4), pure background."""
    rng = np.random.default_rng(seed)
    out = []
    L = len(loci)
    for i in range(npairs):
        u = rng.random()
        l = int(rng.integers(0, L))
        hap = loci[l][int(rng.integers(0, len(loci[l])))]
        if u < frac_locus:
            a, b = sim_pair(hap, rng, rlen, flen)
            kind = "loc%d" % l
        elif u < frac_locus + frac_chimeric and L > 1:
            l2 = int((l + 1 + rng.integers(0, L - 1)) % L)
            hap2 = loci[l2][int(rng.integers(0, len(loci[l2])))]
            a1, b1 = sim_pair(hap, rng, rlen, flen)
            a2, b2 = sim_pair(hap2, rng, rlen, flen)
            c1, c2 = int(rng.integers(30, 121)), int(rng.integers(30, 121))
            a, b = a1[:c1] + a2[c1:], b1[:c2] + b2[c2:]
            kind = "chi%d_%d" % (l, l2)
        elif u < frac_locus + frac_chimeric + frac_border:
            a, b = sim_pair(hap, rng, rlen, flen)
            c1, c2 = int(rng.integers(40, 140)), int(rng.integers(40, 140))
            a, b = a[:c1] + random_read(rng, rlen - c1), b[:c2] + random_read(rng, rlen - c2)
            kind = "bor%d" % l
        elif u < frac_locus + frac_chimeric + frac_border + frac_onemate:
            # one mate keeps only a short locus prefix (passes subfilter at position 0, removed by kfilter), the
            # other a locus prefix of random length: exercises the single-mate accept band
            # 2*cth <= hits < 3*cth-1 of the locus vote (cpp:398, 439-440)
            a, b = sim_pair(hap, rng, rlen, flen)
            c1, c2 = int(rng.integers(40, rlen + 1)), int(rng.integers(25, 61))
            a = a[:c1] + random_read(rng, rlen - c1)
            b = b[:c2] + random_read(rng, rlen - c2)
            if rng.random() < 0.5:
                a, b = b, a
            kind = "one%d" % l
        else:
            a, b = random_read(rng, rlen), random_read(rng, rlen)
            kind = "bg"
        if err:
            a, b = mutate(a, rng, err), mutate(b, rng, err)
        v = rng.random()
        if v < frac_n:
            aa = bytearray(a)
            for p in rng.integers(0, rlen, int(rng.integers(1, 4))):
                aa[int(p)] = ord("N")
            a = bytes(aa)
            if rng.random() < 0.2:
                b = b"N" * rlen
        elif v < frac_n + frac_lower:
            bb = bytearray(b)
            p = int(rng.integers(0, rlen))
            bb[p] = ord(chr(bb[p]).lower())
            b = bytes(bb)
        elif v < frac_n + frac_lower + frac_short:
            a = a[: int(rng.integers(1, 90))]
        out.append((">r%d.%s" % (i, kind), a, b))
    return out


def write_fasta(fn, reads):
    with open(fn, "wb") as f:
        for t, a, b in reads:
            f.write(t.encode() + b"/1\n" + a + b"\n" + t.encode() + b"/2\n" + b + b"\n")
