"""ctypes mirror of include/danbing_b200.h -- plumbing for tests/ and bench.py only.

The product is libdanbing_b200.so (C++/CUDA, danbing-tk_b200/csrc + host); nothing here computes.
Struct layouts must match include/danbing_b200.h field for field.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DBT_LIB") or os.path.join(HERE, "libdanbing_b200.so")   # DBT_LIB: kernel-variant experiments only

DBT_N_COUNTERS = 32
DBT_N_SLOTS = 2
DBT_NCCL_ID_BYTES = 128

F_KAM, F_QC, F_TRACE, F_THREADING, F_CORRECTION, F_ALN, F_ALN_MIN = 1, 2, 4, 8, 16, 32, 64
F_EXTRACT, F_SIM, F_BUBBLES, F_BAIT = 128, 256, 512, 1024
ST_SHORT, ST_SUBFILTERED, ST_KFILTERED, ST_UNASSIGNED, ST_QCFILTERED, ST_ASGN_REMOVED, ST_COUNTED, ST_NOT_THREADED, ST_THREAD_FAILED = range(9)
LAYOUT_ASCII, LAYOUT_PACKED = 0, 1
COUNTER_NAMES = ["nReads", "nShort", "nSubFiltered", "nKmerFiltered", "nBaitFiltered", "nQualFiltered",
                 "nLocusAssignFiltered", "nQCFiltered", "nThreadingReads", "nFeasibleReads", "nAsgnReads",
                 "nhash0", "nhash1", "ninc", "ncandidates", "nrecords",
                 "gen_pairs", "gen_nhash1", "gen_nThreadingReads", "gen_ninc",
                 "thr_walk", "thr_ec", "thr_cls", "thr_refresh",
                 "thr2_pairs", "thr2_mates", "thr2_walk", "thr2_cls", "thr2_ninc", "thr2_refresh", "thr1_fixed", "reserved31"]
MEASUREMENT_COUNTERS = ("nrecords", "gen_pairs", "gen_nhash1", "gen_nThreadingReads", "gen_ninc",
                        "thr_walk", "thr_ec", "thr_cls", "thr_refresh",
                        "thr2_pairs", "thr2_mates", "thr2_walk", "thr2_cls", "thr2_ninc", "thr2_refresh", "thr1_fixed", "reserved31")   # no reference counterpart
MF_KF, MF_HF, MF_BF, MF_QF, MF_AF, MF_RM = 1, 2, 4, 8, 16, 32

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)


class Params(C.Structure):
    _fields_ = [("device", C.c_int32), ("k", C.c_uint32), ("n_filter", C.c_uint32), ("nm_filter", C.c_uint32),
                ("cth", C.c_uint32), ("nm_tr", C.c_uint32), ("max_nt", C.c_uint32), ("thread_cth", C.c_uint32),
                ("max_ncorrection", C.c_uint32), ("flags", C.c_uint32), ("max_read_len", C.c_uint32),
                ("max_batch_pairs", C.c_uint32), ("max_records", C.c_uint32), ("table_load", C.c_float),
                ("bucket_slots", C.c_uint32), ("qth", C.c_uint32), ("reserved", C.c_uint32 * 2)]


class Index(C.Structure):
    _fields_ = [("nloci", C.c_uint64), ("nk_dbi", C.c_uint64), ("dbi_keys", u64p), ("dbi_vals", u32p),
                ("nvv", C.c_uint64), ("vv", u32p), ("tr_n", u64p), ("tr_kmers", u64p), ("fl_n", u64p),
                ("fl_kmers", u64p), ("qc", u8p), ("g_n", u64p), ("g_kmers", u64p), ("g_masks", u8p),
                ("bt_n", u64p), ("bt_kmers", u64p), ("bt_vals", C.POINTER(C.c_uint16)), ("tre_n", u64p), ("tre_kmers", u64p)]


class Batch(C.Structure):
    _fields_ = [("first_pair_index", C.c_uint64), ("npairs", C.c_uint64), ("layout", C.c_int32),
                ("bases", u8p), ("offsets", u64p), ("quals", u8p)]


class PredInput(C.Structure):                      # dbt_pred_input
    _fields_ = [("ns", C.c_uint64), ("nk", C.c_uint64), ("n_tr", C.c_uint64), ("n_ik", C.c_uint64), ("kmc", C.POINTER(u64p)), ("rd", C.POINTER(C.c_float)),
                ("nk_cum", u32p), ("nik_cum", u32p), ("iki", u32p), ("ikmc", u8p)]


class RawInfo(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("nlines", "nrecords", "npairs", "nsingle", "nshort_pairs", "flags", "max_read_len", "max_record_bytes")]


class RawBatch(C.Structure):
    _fields_ = [("first_pair_index", C.c_uint64), ("text", u8p), ("nbytes", C.c_uint64), ("is_fastq", C.c_int32), ("min_read_len", C.c_uint32),
                ("pair_offset", u32p), ("pair_hash", u64p), ("single_offset", u32p), ("single_cap", C.c_uint32)]


RAW_SLOWPATH, RAW_OVERFLOW = 1, 2


class BatchResult(C.Structure):
    _fields_ = [("n_records", C.c_uint64), ("capacity", C.c_uint64), ("records", u8p),
                ("counters", C.c_uint64 * DBT_N_COUNTERS)]


MATE_DTYPE = np.dtype([("si", "<i2"), ("ei", "<i2"), ("si_", "<i2"), ("ei_", "<i2"), ("ti", "<i2"),
                       ("nt", "i1"), ("bs", "i1"), ("flags", "u1"), ("pad", "u1"), ("nk", "<u2")])
RECORD_HEADER_DTYPE = np.dtype([("pair_index", "<u8"), ("dst", "<u4"), ("dst0", "<u4"), ("fc", "<u4"), ("rc", "<u4"),
                                ("nhash0", "<u2"), ("nhash1", "<u2"), ("status", "u1"), ("pad", "u1", 3),
                                ("mate", MATE_DTYPE, 2)])
assert MATE_DTYPE.itemsize == 16 and RECORD_HEADER_DTYPE.itemsize == 64


def record_dtype(as_stride):
    return np.dtype([("h", RECORD_HEADER_DTYPE), ("as", "u1", (2, as_stride))])


ALN_MATE_DTYPE = np.dtype([("alned", "u1"), ("err", "u1"), ("nes", "<u2"), ("ntr", "<u2"), ("nkm", "<u2"), ("kmers_hash", "<u8")])
assert ALN_MATE_DTYPE.itemsize == 16
ES_T = "*=XDI"          # dbt_aln_mate es byte: bits 0-2
ES_G = "_ACGT*"         # bits 3-5 ('_' = none)
TR_C = "*.="


def aln_caps(max_read_len, k, max_ncorrection):
    """(es_cap, tr_cap, aln_stride); must equal dbt_record_es_cap() / dbt_record_aln_stride()."""
    grow = 2 * max_ncorrection + 8
    es_cap = (max_read_len + grow + 7) // 8 * 8
    tr_cap = (max(max_read_len - k + 1, 1) + grow + 7) // 8 * 8
    return es_cap, tr_cap, 16 + es_cap + tr_cap


def aln_record_dtype(es_cap, tr_cap):
    m = np.dtype([("h", ALN_MATE_DTYPE), ("es", "u1", es_cap), ("tr", "u1", tr_cap)])
    return np.dtype([("h", RECORD_HEADER_DTYPE), ("aln", m, 2)])


def fnv1a_u64(kmers):
    h = 0xcbf29ce484222325
    for b in np.asarray(kmers, "<u8").tobytes():
        h = ((h ^ b) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return h


def as_stride_for(max_read_len, k):
    """bytes of one mate's 2-bit annotation; must equal dbt_record_as_stride()."""
    max_nk = max(max_read_len - k + 1, 1)
    return ((max_nk + 3) // 4 + 7) // 8 * 8


def ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def make_index_struct(ix):
    """ix: dict of numpy arrays from rpgg_io.load_index. Keeps references alive on the struct."""
    s = Index()
    s.nloci = int(ix["nloci"])
    s.nk_dbi = len(ix["dbi_keys"])
    s.dbi_keys = ptr(ix["dbi_keys"], u64p)
    s.dbi_vals = ptr(ix["dbi_vals"], u32p)
    s.nvv = len(ix["vv"])
    s.vv = ptr(ix["vv"], u32p) if len(ix["vv"]) else None
    s.tr_n = ptr(ix["tr_n"], u64p)
    s.tr_kmers = ptr(ix["tr_kmers"], u64p)
    s.fl_n = ptr(ix["fl_n"], u64p)
    s.fl_kmers = ptr(ix["fl_kmers"], u64p)
    s.qc = ptr(ix.get("qc"), u8p) if ix.get("qc") is not None else None
    if ix.get("g_n") is not None:                   # PREF.graph.kmers (restored threading)
        s.g_n = ptr(ix["g_n"], u64p)
        s.g_kmers = ptr(ix["g_kmers"], u64p)
        s.g_masks = ptr(ix["g_masks"], u8p)
    if ix.get("bt_n") is not None:                  # PREF.bt.kmdb (`-b`)
        s.bt_n = ptr(ix["bt_n"], u64p)
        s.bt_kmers = ptr(ix["bt_kmers"], u64p)
        s.bt_vals = ix["bt_vals"].ctypes.data_as(C.POINTER(C.c_uint16))
    if ix.get("tre_n") is not None:                 # PREF.tre.kdb (`-bu`)
        s.tre_n = ptr(ix["tre_n"], u64p)
        s.tre_kmers = ptr(ix["tre_kmers"], u64p)
    s._keep = ix
    return s


def default_params(k=21, cth=10, n_filter=4, nm_filter=1, nm_tr=40, flags=F_KAM, max_read_len=150,
                   max_batch_pairs=1 << 20, device=0):
    p = Params()
    p.device = device
    p.k, p.n_filter, p.nm_filter, p.cth, p.nm_tr, p.max_nt = k, n_filter, nm_filter, cth, nm_tr, 2
    p.thread_cth, p.max_ncorrection, p.qth = 100, 4, 20
    p.flags = flags
    p.max_read_len = max_read_len
    p.max_batch_pairs = max_batch_pairs
    return p


_lib = None


def load_library(path=LIB_PATH):
    """Loads libdanbing_b200.so and declares every prototype of include/danbing_b200.h.
    Fails loudly if the library has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise RuntimeError(f"{path} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "or `make -C danbing-tk_b200` (there is no CPU fallback)")
    L = C.CDLL(path)
    ctxp = C.c_void_p
    L.dbt_abi_version.restype = C.c_int
    L.dbt_device_count.restype = C.c_int
    L.dbt_default_params.argtypes = [C.POINTER(Params)]
    L.dbt_create.argtypes = [C.POINTER(ctxp), C.POINTER(Params)]
    L.dbt_destroy.argtypes = [ctxp]
    L.dbt_strerror.argtypes = [C.c_int]
    L.dbt_strerror.restype = C.c_char_p
    L.dbt_last_error.restype = C.c_char_p
    L.dbt_load_index.argtypes = [ctxp, C.POINTER(Index)]
    L.dbt_upload.argtypes = [ctxp, C.POINTER(Batch), C.c_int]
    L.dbt_upload_raw.argtypes = [ctxp, C.POINTER(RawBatch), C.POINTER(RawInfo), C.c_int]
    L.dbt_prepare.argtypes = [ctxp, C.c_uint64]
    L.dbt_graph_merged.argtypes = [ctxp]
    L.dbt_bubble_events.argtypes = [ctxp, C.c_int, u64p, u64p, u32p, C.c_uint64, u64p]
    L.dbt_title_hash.argtypes = [C.c_char_p, C.c_size_t]
    L.dbt_title_hash.restype = C.c_uint64
    L.dbt_raw_times.argtypes = [ctxp, C.c_int, C.POINTER(C.c_float)]
    L.dbt_process.argtypes = [ctxp, C.c_int]
    L.dbt_fetch.argtypes = [ctxp, C.POINTER(BatchResult), C.c_int]
    L.dbt_submit.argtypes = [ctxp, C.POINTER(Batch), C.POINTER(BatchResult), C.c_int]
    L.dbt_wait.argtypes = [ctxp, C.c_int]
    L.dbt_download.argtypes = [ctxp, u64p, u64p, u32p, u64p]
    L.dbt_reset_counts.argtypes = [ctxp]
    for f in (L.dbt_nk_tr, L.dbt_nloci, L.dbt_launch_count, L.dbt_graph_table_bytes):
        f.argtypes = [ctxp]
        f.restype = C.c_uint64
    for f in (L.dbt_record_stride, L.dbt_record_as_stride, L.dbt_record_aln_stride, L.dbt_record_es_cap):
        f.argtypes = [ctxp]
        f.restype = C.c_uint32
    L.dbt_nccl_unique_id.argtypes = [u8p]
    L.dbt_comm_init.argtypes = [ctxp, C.c_int, C.c_int, u8p]
    L.dbt_allreduce.argtypes = [ctxp]
    L.dbt_timer_start.argtypes = [ctxp, C.c_int]
    L.dbt_timer_stop.argtypes = [ctxp, C.c_int, C.POINTER(C.c_float)]
    L.dbt_kernel_times.argtypes = [ctxp, C.c_int, C.POINTER(C.c_float), u32p]
    L.dbt_table_stats.argtypes = [ctxp, u64p]
    L.dbt_batch_stats.argtypes = [ctxp, C.c_int, u64p]
    L.dbt_random_gather_bench.argtypes = [C.c_int, C.c_uint64, C.c_uint32, C.c_uint64, C.c_int, C.POINTER(C.c_float)]
    L.dbt_init_device.argtypes = [C.c_int]
    L.dbt_pred.argtypes = [C.c_int, C.POINTER(PredInput), C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.dbt_noalloc_scratch_needed.argtypes = [C.c_uint32, C.c_uint64, C.c_uint64]
    L.dbt_noalloc_scratch_needed.restype = C.c_size_t
    L.dbt_pack_reads_device_noalloc.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint64, u8p, u64p, u8p]
    L.dbt_bus_bench.argtypes = [C.c_int, C.c_uint64, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.dbt_host_alloc.argtypes = [C.c_size_t]
    L.dbt_host_alloc.restype = C.c_void_p
    L.dbt_host_free.argtypes = [C.c_void_p]
    L.dbt_packed_bytes.argtypes = [C.c_uint32, C.c_uint64]
    L.dbt_packed_bytes.restype = C.c_size_t
    L.dbt_pack_reads.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64, u8p, u64p, u8p, C.c_int]
    _lib = L
    return L


ABI_SYMBOLS = ["dbt_graph_merged", "dbt_bubble_events", "dbt_prepare", "dbt_upload_raw", "dbt_title_hash", "dbt_raw_times", "dbt_abi_version", "dbt_device_count", "dbt_default_params", "dbt_create", "dbt_destroy",
               "dbt_strerror", "dbt_last_error", "dbt_load_index", "dbt_upload", "dbt_process", "dbt_fetch",
               "dbt_submit", "dbt_wait", "dbt_download", "dbt_reset_counts", "dbt_nk_tr", "dbt_nloci",
               "dbt_record_stride", "dbt_record_as_stride", "dbt_record_aln_stride", "dbt_record_es_cap", "dbt_nccl_unique_id", "dbt_comm_init", "dbt_allreduce",
               "dbt_timer_start", "dbt_timer_stop", "dbt_kernel_times", "dbt_launch_count", "dbt_table_stats", "dbt_graph_table_bytes", "dbt_batch_stats",
               "dbt_random_gather_bench", "dbt_init_device", "dbt_pred", "dbt_noalloc_scratch_needed", "dbt_pack_reads_device_noalloc", "dbt_bus_bench", "dbt_host_alloc", "dbt_host_free", "dbt_packed_bytes", "dbt_pack_reads"]


class DbtError(RuntimeError):
    pass


def check(L, rc, what):
    if rc != 0:
        raise DbtError(f"{what}: {L.dbt_strerror(rc).decode()} ({rc}): {L.dbt_last_error().decode()}")


def pinned_array(L, nbytes):
    """uint8 numpy view over cudaHostAlloc'ed memory (freed when the returned owner is collected)."""
    p = L.dbt_host_alloc(nbytes)
    if not p:
        raise MemoryError("dbt_host_alloc")
    buf = (C.c_uint8 * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=np.uint8)

    class _Owner:
        def __init__(self, p):
            self.p = p

        def __del__(self):
            if self.p:
                L.dbt_host_free(self.p)
                self.p = None
    return arr, _Owner(p)


class Context:
    """Thin RAII wrapper over dbt_ctx for tests and bench.py."""

    def __init__(self, params, index_struct=None):
        self.L = load_library()
        self.params = params
        self.h = C.c_void_p()
        check(self.L, self.L.dbt_create(C.byref(self.h), C.byref(params)), "dbt_create")
        if index_struct is not None:
            self.load_index(index_struct)

    def load_index(self, index_struct):
        check(self.L, self.L.dbt_load_index(self.h, C.byref(index_struct)), "dbt_load_index")
        self.nk_tr = self.L.dbt_nk_tr(self.h)
        self.nloci = self.L.dbt_nloci(self.h)
        self.stride = self.L.dbt_record_stride(self.h)
        self.as_stride = self.L.dbt_record_as_stride(self.h)
        self.aln_stride = self.L.dbt_record_aln_stride(self.h)
        self.es_cap = self.L.dbt_record_es_cap(self.h)
        self.threading = self.stride == 64 + 2 * self.aln_stride      # restored threading is active (graph loaded)

    def close(self):
        if self.h:
            self.L.dbt_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def make_batch(self, bases, offsets, first_pair_index=0, layout=LAYOUT_ASCII, npairs=None):
        b = Batch()
        b.first_pair_index = first_pair_index
        b.npairs = (len(offsets) - 1) // 2 if npairs is None else npairs
        b.layout = layout
        b.bases = ptr(bases, u8p)
        b.offsets = ptr(offsets, u64p) if offsets is not None else None
        b._keep = (bases, offsets)
        return b

    def run(self, bases, offsets, first_pair_index=0, capacity=None, slot=0):
        """submit + wait; returns (records ndarray, counters ndarray)."""
        b = self.make_batch(bases, offsets, first_pair_index)
        cap = int(b.npairs if capacity is None else capacity)
        rec = np.zeros(max(cap, 1) * self.stride, dtype=np.uint8)
        r = BatchResult()
        r.capacity = cap
        r.records = ptr(rec, u8p)
        check(self.L, self.L.dbt_submit(self.h, C.byref(b), C.byref(r), slot), "dbt_submit")
        check(self.L, self.L.dbt_wait(self.h, slot), "dbt_wait")
        n = int(r.n_records)
        dt = aln_record_dtype(self.es_cap, self.aln_stride - 16 - self.es_cap) if self.threading else record_dtype(self.as_stride)
        recs = rec[: n * self.stride].view(dt)
        return recs, np.array(list(r.counters), dtype=np.uint64)

    def run_raw(self, text, is_fastq, min_read_len, first_pair_index=0, slot=0, single_cap=1 << 16):
        """dbt_upload_raw + process + fetch + wait on raw FASTA/FASTQ text (bytes).
        -> (info dict, pair_offset, pair_hash, single_offset, records, counters); records None when info['flags'] != 0."""
        t = np.frombuffer(text, dtype=np.uint8)
        P = int(self.params.max_batch_pairs)
        po, ph, so = np.zeros(P, "<u4"), np.zeros(P, "<u8"), np.zeros(single_cap, "<u4")
        b = RawBatch()
        b.first_pair_index, b.text, b.nbytes, b.is_fastq, b.min_read_len = first_pair_index, ptr(t, u8p), len(t), int(is_fastq), min_read_len
        b.pair_offset, b.pair_hash, b.single_offset, b.single_cap = ptr(po, u32p), ptr(ph, u64p), ptr(so, u32p), single_cap
        info = RawInfo()
        check(self.L, self.L.dbt_upload_raw(self.h, C.byref(b), C.byref(info), slot), "dbt_upload_raw")
        d = {n: int(getattr(info, n)) for n, _ in RawInfo._fields_}
        if d["flags"]:
            return d, None, None, None, None, None
        n = d["npairs"]
        rec = np.zeros(max(n, 1) * self.stride, dtype=np.uint8)
        r = BatchResult()
        r.capacity = n
        r.records = ptr(rec, u8p)
        check(self.L, self.L.dbt_process(self.h, slot), "dbt_process")
        check(self.L, self.L.dbt_fetch(self.h, C.byref(r), slot), "dbt_fetch")
        check(self.L, self.L.dbt_wait(self.h, slot), "dbt_wait")
        dt = aln_record_dtype(self.es_cap, self.aln_stride - 16 - self.es_cap) if self.threading else record_dtype(self.as_stride)
        recs = rec[: int(r.n_records) * self.stride].view(dt)
        return d, po[:n].copy(), ph[:n].copy(), so[:d["nsingle"]].copy(), recs, np.array(list(r.counters), dtype=np.uint64)

    def download(self):
        trkmc = np.zeros(max(self.nk_tr, 1), dtype=np.uint64)
        kmc = np.zeros(max(self.nloci, 1), dtype=np.uint64)
        nmap = np.zeros(max(self.nloci, 1), dtype=np.uint32)
        cnt = np.zeros(DBT_N_COUNTERS, dtype=np.uint64)
        check(self.L, self.L.dbt_download(self.h, ptr(trkmc, u64p), ptr(kmc, u64p), ptr(nmap, u32p), ptr(cnt, u64p)),
              "dbt_download")
        return trkmc[: self.nk_tr], kmc[: self.nloci], nmap[: self.nloci], cnt

    def reset(self):
        check(self.L, self.L.dbt_reset_counts(self.h), "dbt_reset_counts")
