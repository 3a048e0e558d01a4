"""ctypes access to the CPU oracle (oracle/liboracle.so) and the compiled reference (oracle/_ref).

TEST INFRASTRUCTURE: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg only.
"""
import ctypes as C
import importlib.util
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
PKG_DIR = os.path.join(ROOT, "danbing-tk_b200")


def load_pkg():
    """The package directory name has a hyphen (it mirrors the reference's name); import it by path."""
    name = "danbing_tk_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(PKG_DIR, "__init__.py"),
                                                  submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


pkg = load_pkg()
abi = pkg.abi
rpgg_io = pkg.rpgg_io


def ref_bin(name):
    p = os.path.join(REF_DIR, name)
    return p if os.path.exists(p) else None


def have_ref():
    return all(ref_bin(b) for b in ("danbing-tk", "ktools", "fa2kmers", "ref_harness"))


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "port"], check=True)


_lib = None


def oracle_lib():
    global _lib
    if _lib is None:
        so = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(so):
            build_oracle()
        L = C.CDLL(so)
        L.dbo_index_new.argtypes = [C.POINTER(abi.Index)]
        L.dbo_index_new.restype = C.c_void_p
        L.dbo_index_free.argtypes = [C.c_void_p]
        L.dbo_run.argtypes = [C.c_void_p, C.POINTER(abi.Params), C.c_uint64, C.c_uint64, abi.u8p, abi.u64p,
                              abi.u8p, C.c_uint32, C.c_uint32, abi.u64p, abi.u64p, abi.u32p, abi.u64p]
        L.dbo_revcomp.argtypes = [C.c_uint64, C.c_uint32]
        L.dbo_revcomp.restype = C.c_uint64
        L.dbo_read2kmers_edges.argtypes = [abi.u8p, C.c_uint32, C.c_uint32, abi.u64p, abi.u64p]
        L.dbo_read2kmers_edges.restype = C.c_uint32
        L.dbo_introsort_by_key.argtypes = [abi.u64p, C.c_uint32, abi.u64p]
        _lib = L
    return _lib


class Oracle:
    def __init__(self, ix):
        self.L = oracle_lib()
        self.ix = ix
        self.s = abi.make_index_struct(ix)
        self.h = self.L.dbo_index_new(C.byref(self.s))
        assert self.h
        self.nloci = int(ix["nloci"])
        self.nk_tr = len(ix["tr_kmers"])

    def __del__(self):
        if getattr(self, "h", None):
            self.L.dbo_index_free(self.h)
            self.h = None

    def run(self, params, bases, offsets, first_pair_index=0, acc=None):
        """Returns (records, counters, trkmc, kmc, nmapread). acc = previous (trkmc,kmc,nmapread,counters) to accumulate."""
        npairs = (len(offsets) - 1) // 2
        max_len = int(np.max(np.diff(offsets.astype(np.int64)))) if npairs else 1
        as_stride = abi.as_stride_for(max(max_len, params.max_read_len), params.k)
        stride = 64 + 2 * as_stride
        threading = bool(params.flags & abi.F_THREADING) and self.ix.get("g_n") is not None
        if threading:                         # restored threading: header + two dbt_aln_mate blobs
            assert max_len <= params.max_read_len
            es_cap, tr_cap, aln_stride = abi.aln_caps(params.max_read_len, params.k, params.max_ncorrection)
            stride = 64 + 2 * aln_stride
        rec = np.zeros(max(npairs, 1) * stride, np.uint8)
        if acc is None:
            trkmc = np.zeros(max(self.nk_tr, 1), np.uint64)
            kmc = np.zeros(max(self.nloci, 1), np.uint64)
            nmap = np.zeros(max(self.nloci, 1), np.uint32)
            cnt = np.zeros(abi.DBT_N_COUNTERS, np.uint64)
        else:
            trkmc, kmc, nmap, cnt = acc
        self.L.dbo_run(self.h, C.byref(params), first_pair_index, npairs, abi.ptr(bases, abi.u8p),
                       abi.ptr(offsets, abi.u64p), abi.ptr(rec, abi.u8p), stride, as_stride,
                       abi.ptr(trkmc, abi.u64p), abi.ptr(kmc, abi.u64p), abi.ptr(nmap, abi.u32p), abi.ptr(cnt, abi.u64p))
        recs = rec[: npairs * stride].view(abi.aln_record_dtype(es_cap, tr_cap) if threading else abi.record_dtype(as_stride))
        return recs, cnt, trkmc[: self.nk_tr], kmc[: self.nloci], nmap[: self.nloci]


def run_ref_harness(pref, seqs, outpref, k, cth, nf, nmf, nmtr, okam=True, qcfile=None):
    """Runs oracle/_ref/ref_harness; returns (trace lines, trkmc, summary text, counters dict)."""
    tsv = outpref + ".pairs.tsv"
    rpgg_io.write_pairs_tsv(tsv, seqs)
    cmd = [ref_bin("ref_harness"), pref, tsv, outpref, str(k), str(cth), str(nf), str(nmf), str(nmtr), str(int(okam))]
    if qcfile:
        cmd.append(qcfile)
    subprocess.run(cmd, check=True, stderr=subprocess.DEVNULL)
    trace = open(outpref + ".trace.tsv").read().splitlines()
    trkmc = rpgg_io.read_ar(outpref + ".trkmc.ar")
    summary = open(outpref + ".tr.summary.txt").read()
    counters = dict((l.split("\t")[0], int(l.split("\t")[1])) for l in open(outpref + ".counters.txt").read().splitlines())
    return trace, trkmc, summary, counters


def run_ref_binary(pref, fasta, outpref, k=21, cth=10, kf=(4, 1), c=40, p=1, named=False, extra=()):
    """Runs the unmodified reference `danbing-tk`; returns (kam text, stderr text)."""
    cmd = [ref_bin("danbing-tk"), "-k", str(k), "-kf", str(kf[0]), str(kf[1]), "-cth", str(cth), "-c", str(c), "-p", str(p)]
    cmd += list(extra) + ["-qs", pref, "-fa", fasta, "-on" if named else "-o", outpref]
    r = subprocess.run(cmd, check=True, capture_output=True, text=True)
    return r.stdout, r.stderr
