"""Inputs and a float32 restatement of `danbing-tk-pred` (reference src/pred.cpp, src/pred.h) for the tests.

TEST INFRASTRUCTURE.  pred_model() follows the reference operation by operation in IEEE float32, including the order in which
Eigen 3.4 (vendored in the reference tree, SSE2 packets of 4 floats) adds things up:
  * norm_rd (pred.h:166-171):     gt[s, k] = float(kmc[s][k]) / rd[s]
  * bias_correction (pred.h:173-195), per locus with k-mers [si, ei) and invariant k-mers [isi, iei):
        B[s, j]  = gt[s, iki[j]] / ikmc[j]
        bias[s]  = (B[s,0] + B[s,1] + ... left to right) / float(n_ik)          rowwise().mean() over an indexed view: scalar loop
        bias    /= mean(bias)                                                    full reduction of a contiguous array: two packet
                                                                                 accumulators, then (a0+a2)+(a1+a3), then the tail
        gt[s, si:ei] /= bias[s];   Bias[s, locus] = bias[s];   loci without k-mers or without invariant k-mers are skipped
tests/test_pred.py pins this model bit for bit against the compiled reference (oracle/_ref/danbing-tk-pred)."""
import os
import struct

import numpy as np

f32 = np.float32


def make_pred_case(d, ns=12, n_tr=40, seed=1, kmax=60, skip_some=False):
    """writes ns OUT.trkmc.ar files, trkmers.meta.txt and ikmer.meta under d; returns the arrays"""
    rng = np.random.default_rng(seed)
    nk_l = rng.integers(10, kmax, n_tr)
    nik_l = np.array([rng.integers(1, max(2, n // 3)) for n in nk_l])
    if skip_some:                                     # loci the reference skips: no invariant k-mers
        nik_l[rng.integers(0, n_tr, max(1, n_tr // 8))] = 0
    nk_cum = np.cumsum(nk_l).astype(np.uint32)
    nik_cum = np.cumsum(nik_l).astype(np.uint32)
    nk, nik = int(nk_cum[-1]), int(nik_cum[-1])
    iki = np.zeros(nik, np.uint32)
    ikmc = np.zeros(nik, np.uint8)
    for t in range(n_tr):
        si = int(nk_cum[t - 1]) if t else 0
        isi = int(nik_cum[t - 1]) if t else 0
        n = int(nik_l[t])
        iki[isi:isi + n] = si + np.sort(rng.choice(int(nk_l[t]), n, replace=False))
        ikmc[isi:isi + n] = rng.integers(1, 5, n)
    rd = rng.uniform(8.0, 45.0, ns).round(3)
    kmc = []
    fns = []
    for s in range(ns):
        lam = np.repeat(rng.uniform(0.7, 1.3, n_tr), nk_l) * rd[s] * rng.integers(1, 4, nk)
        v = rng.poisson(lam).astype(np.uint64)
        v[iki] = np.maximum(v[iki], 1)               # an all-zero invariant set would make bias 0 -> inf / nan in both arms
        kmc.append(v)
        fn = os.path.join(d, f"s{s}.trkmc.ar")
        with open(fn, "wb") as f:
            f.write(struct.pack("<Q", nk))
            f.write(v.tobytes())
        fns.append(fn)
    meta = os.path.join(d, "trkmers.meta.txt")
    with open(meta, "w") as f:
        for fn, r in zip(fns, rd):
            f.write(f"{fn}\t{r}\n")
    ikfn = os.path.join(d, "ikmer.meta")
    with open(ikfn, "wb") as f:                       # read_ikmer, pred.h:69-123
        f.write(struct.pack("<QQQ", nk, nik, n_tr))
        f.write(nk_cum.tobytes())
        f.write(nik_cum.tobytes())
        for a, b in zip(iki, ikmc):
            f.write(struct.pack("<IB", int(a), int(b)))
    rdf = np.array([f32(float(f"{r}")) for r in rd], f32)      # stof of the printed text
    return dict(ns=ns, nk=nk, n_tr=n_tr, nik=nik, kmc=kmc, rd=rdf, nk_cum=nk_cum, nik_cum=nik_cum, iki=iki, ikmc=ikmc,
                meta=meta, ikfn=ikfn, dir=d)


def eigen_sum_f32(a):
    """Eigen 3.4 redux (LinearVectorizedTraversal, NoUnrolling) of a contiguous, 16-byte-aligned float array with SSE packets"""
    a = np.asarray(a, f32)
    n = len(a)
    P = 4
    n2 = (n // (2 * P)) * (2 * P)
    n1 = (n // P) * P
    if n1 == 0:
        r = a[0]
        for i in range(1, n):
            r = f32(r + a[i])
        return r
    p0 = a[0:P].copy()
    if n1 > P:
        p1 = a[P:2 * P].copy()
        for i in range(2 * P, n2, 2 * P):
            p0 = (p0 + a[i:i + P]).astype(f32)
            p1 = (p1 + a[i + P:i + 2 * P]).astype(f32)
        p0 = (p0 + p1).astype(f32)
        if n1 > n2:
            p0 = (p0 + a[n2:n2 + P]).astype(f32)
    r = f32(f32(p0[0] + p0[2]) + f32(p0[1] + p0[3]))
    for i in range(n1, n):
        r = f32(r + a[i])
    return r


def pred_model(c):
    """-> raw (ns, nk), corrected (ns, nk), bias (ns, n_tr) float32; bias columns of skipped loci are NaN"""
    ns, nk, n_tr = c["ns"], c["nk"], c["n_tr"]
    raw = np.empty((ns, nk), f32)
    for s in range(ns):
        raw[s] = (c["kmc"][s].astype(f32) / c["rd"][s]).astype(f32)
    cor = raw.copy()
    bias_out = np.full((ns, n_tr), np.nan, f32)
    ikf = c["ikmc"].astype(f32)
    for t in range(n_tr):
        si = int(c["nk_cum"][t - 1]) if t else 0
        ei = int(c["nk_cum"][t])
        isi = int(c["nik_cum"][t - 1]) if t else 0
        iei = int(c["nik_cum"][t])
        if si == ei or isi == iei:
            continue
        cols = c["iki"][isi:iei].astype(np.int64)
        B = (cor[:, cols] / ikf[isi:iei][None, :]).astype(f32)
        acc = B[:, 0].copy()
        for j in range(1, B.shape[1]):
            acc = (acc + B[:, j]).astype(f32)
        bias = (acc / f32(iei - isi)).astype(f32)
        m = f32(eigen_sum_f32(bias) / f32(ns))
        bias = (bias / m).astype(f32)
        cor[:, si:ei] = (cor[:, si:ei] / bias[:, None]).astype(f32)
        bias_out[:, t] = bias
    return raw, cor, bias_out


def read_matrix(fn):
    """save_matrix (pred.h:197-210): u32 nrow, u32 ncol, column-major float32"""
    b = open(fn, "rb").read()
    nrow, ncol = struct.unpack("<II", b[:8])
    return np.frombuffer(b, f32, nrow * ncol, 8).reshape(ncol, nrow).T


def read_bias_tsv(fn):
    return np.array([[float(x) for x in line.split("\t")] for line in open(fn).read().split("\n") if line], np.float64)
