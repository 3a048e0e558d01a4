#!/usr/bin/env python3
"""bench.py -- read pairs/sec of the `danbing-tk align` hot path (k=21) on N B200s.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line (rank 0).
  step      = one pass of the hot path (subfilter -> compaction -> candidate kernel) over one batch of
              synthetic read pairs; N>1: every rank runs the same-sized batch of ITS OWN reads
              (weak scaling, RPGG replicated) and a step ends with the NCCL sum-reduce of the counts.
  value     = read pairs/s, whole job, packed reads already resident in HBM (CUDA events, max over ranks)
  e2e       = same metric through the C ABI with HOST buffers: pinned ASCII reads -> dbt_submit (H2D +
              device 2-bit pack + kernels) -> dbt_wait (D2H of counters + kam/aln records), double-buffered.
              --e2e-input both (default) also times the host-packed and the hybrid input path from the same
              ASCII reads (dbt_pack_reads inside the timed region) and reports the fastest, the rest in other_paths
  roofline  = dominant kernel's algorithmic HBM bytes / its CUDA-event time vs MEASURED_PEAKS.json
  cpu_baseline = the UNMODIFIED reference (oracle/_ref/danbing-tk -p nproc) on a bounded sample of the
              same workload on this box's host cores, outputs diffed against the GPU's (bit-exact gate)
`--impl reference` times only that CPU reference arm.
Workload = BASELINE.json configs[1]: synthetic 1k-locus RPGG, 10M simulated 150 bp pairs (5% locus
reads, 95% background), flags -k 21 -gc 85 3 -kf 4 1 -cth 45 -ae.
Two modes, both measured in the default run:
  thread (HEADLINE)  restored graph threading (SURVEY.md §8(c) decision D1): locus reads carry 0.2 %
              substitutions + 0.01 % indels (S2), every pair that passes the locus vote is threaded through
              the locus graph with error correction, counts come from the edited k-mer vectors, `-ae`
              alignment records come back.  CPU arm = the reference's own isThreadFeasible & helpers
              compiled from /root/reference (oracle/_ref/ref_harness; the commented-out caller
              cpp:2072-2088 restored as glue), one process per host core on shards of the sample.
  asgn (block "asgn")  what the UNMODIFIED reference binary executes at HEAD for these flags minus -gc
              (threading has no caller there): error-free reads, kam records.  CPU arm =
              oracle/_ref/danbing-tk -p nproc.
"""
import argparse
import ctypes as C
import json
import re
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_abi import REF_DIR, abi, pkg, ref_bin, rpgg_io  # noqa: E402  (oracle_abi: only the cpu_baseline leg runs oracle/_ref)

synth = pkg.synth
RLEN, FLEN, K, CTH = 150, 500, 21, 45
CORR = 1                                   # 1: `-gc 85 3` (error-correcting threading), 0: `-g 85` (exact threading)


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------
def build_workload(nloci, seed, workdir):
    """S2-shaped RPGG: nloci loci x 8 haplotypes, 700 bp flanks, 5% sharing a 200 bp segment."""
    t = time.time()
    loci = synth.make_loci(nloci, seed=seed, nhap=8, flank=700, share_frac=0.05, share_len=200)
    pref = os.path.join(workdir, "pan")
    nproc = len(os.sched_getaffinity(0))
    ix = synth.build_rpgg(loci, pref, k=K, fsi=700, fso=700, write_text=True, nproc=nproc)
    ix.update(synth.build_graph(loci, K, pref, nproc=nproc))  # PREF.graph.kmers (fa2kmers -g) for the threading mode
    haps = [h for hs in loci for h in hs]
    hap_off = np.zeros(len(haps) + 1, np.uint64)
    hap_off[1:] = np.cumsum([len(h) for h in haps])
    pool = np.frombuffer(b"".join(haps), np.uint8).copy()
    log(f"[bench] RPGG: {nloci} loci, {len(ix['dbi_keys'])} index k-mers, {len(ix['tr_kmers'])} TR k-mers, nvv={len(ix['vv'])}, {len(ix['g_kmers'])} graph nodes ({time.time()-t:.1f}s)")
    return ix, pref, pool, hap_off


_synth_lib = None


SUB_PPM, INDEL_PPM = 2000, 100          # S2 error model of the threading runs: 0.2 % substitutions, 0.01 % indels


def synth_reads(device, pool, hap_off, first_pair, npairs, seed, frac_locus, out, sub_ppm=0, indel_ppm=0):
    global _synth_lib
    if _synth_lib is None:
        _synth_lib = C.CDLL(os.path.join(ROOT, "danbing-tk_b200", "libdbt_synth.so"))
        _synth_lib.dbt_synth_reads2.argtypes = [C.c_int, abi.u8p, abi.u64p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64,
                                                C.c_double, C.c_uint32, C.c_uint32, abi.u8p, C.c_uint32, C.c_uint32]
    rc = _synth_lib.dbt_synth_reads2(device, abi.ptr(pool, abi.u8p), abi.ptr(hap_off, abi.u64p), len(hap_off) - 1, first_pair,
                                     npairs, seed, frac_locus, RLEN, FLEN, abi.ptr(out, abi.u8p), sub_ppm, indel_ppm)
    if rc:
        raise RuntimeError(f"dbt_synth_reads failed: {rc}")


def synth_reads_host(pool, hap_off, first_pair, npairs, seed, frac_locus, out, sub_ppm=0, indel_ppm=0):
    """The same counter-based generator on the host cores (libdbt_synth.so, no CUDA call): used by the reference arm so that
    nothing of the product and no GPU code runs in it.  Byte-identical to synth_reads."""
    lib = C.CDLL(os.path.join(ROOT, "danbing-tk_b200", "libdbt_synth.so"))
    lib.dbt_synth_reads_host.argtypes = [abi.u8p, abi.u64p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64, C.c_double,
                                         C.c_uint32, C.c_uint32, abi.u8p, C.c_uint32, C.c_uint32, C.c_int]
    rc = lib.dbt_synth_reads_host(abi.ptr(pool, abi.u8p), abi.ptr(hap_off, abi.u64p), len(hap_off) - 1, first_pair, npairs, seed,
                                  frac_locus, RLEN, FLEN, abi.ptr(out, abi.u8p), sub_ppm, indel_ppm, 0)
    if rc:
        raise RuntimeError(f"dbt_synth_reads_host failed: {rc}")


def write_fasta_sample(fn, ascii_reads, npairs):
    """pairs [0, npairs) of the batch as interleaved FASTA in ARRIVAL order (/1 then /2), fixed-width names."""
    a = ascii_reads[: npairs * 2 * RLEN].reshape(npairs, 2, RLEN)
    idw = 9
    rec = np.empty((npairs, 2, 1 + 1 + idw + 2 + 1 + RLEN + 1), np.uint8)
    ids = np.char.zfill(np.arange(npairs).astype(str), idw)
    idb = np.frombuffer("".join(ids.tolist()).encode(), np.uint8).reshape(npairs, idw)
    for m, tag in ((0, b"1"), (1, b"2")):
        rec[:, m, 0] = ord(">")
        rec[:, m, 1] = ord("r")
        rec[:, m, 2:2 + idw] = idb
        rec[:, m, 2 + idw] = ord("/")
        rec[:, m, 3 + idw] = tag[0]
        rec[:, m, 4 + idw] = ord("\n")
        rec[:, m, 5 + idw:5 + idw + RLEN] = a[:, 1 - m]          # batch order is (late "/2", early "/1")
        rec[:, m, 5 + idw + RLEN] = ord("\n")
    rec.tofile(fn)


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.samples, self.stop_flag, self.device = [], False, device
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.device)],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.samples.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        self.t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        reasons = [n for i, n in ((3, "hw_slowdown"), (4, "hw_thermal_slowdown"), (5, "sw_thermal_slowdown"), (6, "sw_power_cap"))
                   if any(s[i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# reference arm (CPU)
# ------------------------------------------------------------------------------------------------
def run_reference(pref, fasta, outpref, nproc):
    """wall time of the reference's query phase = full run minus a zero-read run (index load)."""
    exe = ref_bin("danbing-tk")
    base = [exe, "-k", str(K), "-kf", "4", "1", "-cth", str(CTH), "-ae", "-p", str(nproc), "-qs", pref]
    empty = outpref + ".empty.fa"
    open(empty, "w").close()
    t0 = time.perf_counter()
    subprocess.run(base + ["-fa", empty, "-o", outpref + ".e"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    t_load = time.perf_counter() - t0
    t0 = time.perf_counter()
    subprocess.run(base + ["-fa", fasta, "-o", outpref], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    t_all = time.perf_counter() - t0
    return max(t_all - t_load, 1e-9), t_all, t_load


def write_pairs_tsv_sample(fn, ascii_reads, p0, p1):
    """pairs [p0, p1) of the batch as `seq1<TAB>seq2` lines in the reference's internal mate order (= batch order)."""
    a = ascii_reads[p0 * 2 * RLEN: p1 * 2 * RLEN].reshape(p1 - p0, 2, RLEN)
    rec = np.empty((p1 - p0, 2 * RLEN + 2), np.uint8)
    rec[:, :RLEN] = a[:, 0]
    rec[:, RLEN] = ord("\t")
    rec[:, RLEN + 1: 2 * RLEN + 1] = a[:, 1]
    rec[:, 2 * RLEN + 1] = ord("\n")
    rec.tofile(fn)


def run_reference_thread(pref, ascii_reads, n, workdir, nproc):
    """The reference's threading functions (oracle/_ref/ref_harness, thread=85 corr=1 maxnc=3 aln=2, no trace dump):
    one process per core, each on a contiguous shard of the first n pairs.  -> (query s, wall s, load s, summed trkmc)."""
    exe = ref_bin("ref_harness")
    tail = [str(K), str(CTH), "4", "1", "40", "0", "thread=85", f"corr={CORR}", "maxnc=3", "aln=2", "trace=0"]
    nproc = max(1, min(nproc, n // 1000 or 1))
    bounds = [n * i // nproc for i in range(nproc + 1)]
    empty = os.path.join(workdir, "thr.empty.tsv")
    open(empty, "w").close()
    for i in range(nproc):
        write_pairs_tsv_sample(os.path.join(workdir, f"thr.{i}.tsv"), ascii_reads, bounds[i], bounds[i + 1])

    def launch(tsvs):
        t0 = time.perf_counter()
        ps = [subprocess.Popen([exe, pref, tsv, tsv + ".out"] + tail, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for tsv in tsvs]
        for q in ps:
            if q.wait():
                raise RuntimeError("ref_harness failed")
        return time.perf_counter() - t0
    # index load phase: the same nproc processes on empty inputs (they contend for memory bandwidth the same way)
    empties = [f"{empty}.{i}" for i in range(nproc)]
    for e in empties:
        open(e, "w").close()
    t_load = launch(empties)
    t_all = launch([os.path.join(workdir, f"thr.{i}.tsv") for i in range(nproc)])
    tot = None
    for i in range(nproc):
        c = rpgg_io.read_ar(os.path.join(workdir, f"thr.{i}.tsv.out.trkmc.ar"))
        tot = c.astype(np.uint64) if tot is None else tot + c
    return max(t_all - t_load, 1e-9), t_all, t_load, tot, nproc


def peaks():
    fn = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(fn):
        return float(json.load(open(fn))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
METRIC = "read pairs/sec (k=21, -ae)"
def flags_txt(mode):
    if mode == "thread":
        return f"align -k {K} {'-gc 85 3' if CORR else '-g 85'} -kf 4 1 -cth 45 -ae (restored graph threading, SURVEY §8(c) D1)"
    return f"align -k {K} -kf 4 1 -cth 45 -ae (asgn path: what the unmodified reference runs at HEAD, SURVEY F2)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="both", choices=["both", "thread", "asgn"], help="headline = thread; 'both' adds the asgn block")
    ap.add_argument("--pairs", type=int, default=10_000_000, help="read pairs per GPU per step")
    ap.add_argument("--loci", type=int, default=1000)
    ap.add_argument("--k", type=int, default=21, help="k-mer size (configs[3]: 17 / 25); the metric is quoted at the default 21")
    ap.add_argument("--exact", action="store_true", help="threading mode with `-g 85` (no error correction, configs[4]) instead of `-gc 85 3`")
    ap.add_argument("--frac-locus", type=float, default=0.05)
    ap.add_argument("--cpu-sample", type=int, default=2_000_000, help="pairs given to the CPU reference")
    ap.add_argument("--sub-batch", type=int, default=2_500_000, help="pairs per dbt_submit in the e2e loop")
    ap.add_argument("--e2e-input", choices=("ascii", "hostpack", "hybrid", "both"), default="both",
                    help="e2e input path: ASCII H2D + device packer, host packer + packed H2D, hybrid (host packs every --hybrid-every-th "
                         "sub-batch while PCIe carries ASCII for the others), or `both` = measure all three and report the fastest")
    ap.add_argument("--hybrid-every", type=int, default=3)
    ap.add_argument("--pack-threads", type=int, default=0, help="host packer threads (0 = CPUs in this process's affinity mask)")
    ap.add_argument("--bucket-slots", type=int, default=0)
    ap.add_argument("--table-load", type=float, default=0.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--hbm-loci", type=int, default=20000, help="second, HBM-resident workload (index and graph tables >> L2) measured at N=1 "
                    "as block `hbm_resident`; 0 = off")
    ap.add_argument("--hbm-pairs", type=int, default=2_000_000)
    ap.add_argument("--hbm-cpu-sample", type=int, default=200_000)
    ap.add_argument("--file-pairs", type=int, default=8_000_000, help="block `e2e_file` (N=1): FASTA of this many pairs in /dev/shm -> the drop-in CLI -> "
                    "outputs in /dev/shm, beside the reference binary on the same file; 0 = off")
    ap.add_argument("--pred-samples", type=int, default=48, help="block `pred` (N=1): samples given to danbing-tk-pred (SURVEY §8(f) row 4); 0 = skip")
    ap.add_argument("--pred-kmers", type=int, default=2_000_000)
    ap.add_argument("--rrand", action=argparse.BooleanOptionalAction, default=True,
                    help="also measure the random-sector roofline R_rand (32 B / 64 B gathers over an 8 GiB table, < 1 s) and report the "
                         "probe rate of the dominant kernel as a fraction of it")
    a = ap.parse_args()
    a.steps, a.warmup = max(a.steps, 1), max(a.warmup, 0)
    global K, CORR
    K, CORR = a.k, 0 if a.exact else 1

    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # keep stdout to the one JSON line (NCCL prints its version banner there)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1 and a.impl != "reference":
        import torch.distributed as dist           # plumbing only: rendezvous, barrier, max over ranks
        dist.init_process_group("gloo", rank=rank, world_size=world)

    workdir = tempfile.mkdtemp(prefix="dbt_bench_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        if a.impl == "reference":
            if rank != 0:
                return
            return reference_arm(a, workdir)
        ix, pref, pool, hap_off = build_workload(a.loci, 20261104, workdir)
        head = "asgn" if a.mode == "asgn" else "thread"
        out = b200_mode(head, a, ix, pref, pool, hap_off, workdir, rank, world, local, dist)
        if a.mode == "both":
            sec = b200_mode("asgn", a, ix, pref, pool, hap_off, workdir, rank, world, local, dist)
            if rank == 0:
                out["asgn"] = {k: sec[k] for k in ("value", "unit", "ms_per_step", "e2e", "gpu_launches", "roofline", "cpu_baseline",
                                                   "config", "counters_per_step_rank0")}
        if rank == 0 and world == 1 and a.hbm_loci > 0 and a.hbm_loci != a.loci and a.mode != "asgn":
            out["hbm_resident"] = hbm_resident_block(a, workdir)
        if rank == 0 and world == 1 and a.file_pairs > 0 and a.loci == 1000:
            out["e2e_file"] = e2e_file_block(a, pref, pool, hap_off, workdir, local)
        if rank == 0 and world == 1 and a.pred_samples > 0:
            out["pred"] = pred_block(a, workdir, local)
        if rank == 0:
            print(json.dumps(out))
    finally:
        shutil.rmtree(workdir, ignore_errors=True)
        if dist is not None:
            dist.destroy_process_group()


def pred_block(a, workdir, device):
    """SURVEY §8(f) row 4, `danbing-tk-pred`: --pred-samples count arrays (synthetic OUT.trkmc.ar of --pred-kmers k-mers) through
    dbt_pred (kernel times by CUDA events, algorithmic bytes vs the HBM peak), the drop-in CLI and the unmodified reference binary
    (single-threaded Eigen) on the same files in /dev/shm; the three output files are compared byte for byte."""
    L = abi.load_library()
    ns, nk = a.pred_samples, a.pred_kmers
    rng = np.random.default_rng(77)
    n_tr = max(10, nk // 180)                                 # ~180 k-mers per locus, a fifth of them invariant
    cuts = np.sort(rng.choice(np.arange(1, nk), n_tr - 1, replace=False))
    nk_cum = np.append(cuts, nk).astype(np.uint32)
    starts = np.concatenate([[0], cuts]).astype(np.int64)
    inv = np.flatnonzero(rng.random(nk) < 0.2).astype(np.uint32)
    first = starts[np.searchsorted(nk_cum, starts, side="right") == np.arange(n_tr)]   # every locus keeps at least its first k-mer
    inv = np.union1d(inv, first.astype(np.uint32)).astype(np.uint32)
    nik_cum = np.searchsorted(inv, nk_cum, side="left").astype(np.uint32)
    ikmc = rng.integers(1, 4, len(inv)).astype(np.uint8)
    rd = rng.uniform(20, 40, ns).round(2)
    d = os.path.join(workdir, "pred")
    os.makedirs(d, exist_ok=True)
    kmc, fns = [], []
    for s in range(ns):
        v = (rng.poisson(rd[s], nk) + 1).astype(np.uint64)
        kmc.append(v)
        fn = os.path.join(d, f"s{s}.trkmc.ar")
        with open(fn, "wb") as f:
            f.write(np.uint64(nk).tobytes())
            f.write(v.tobytes())
        fns.append(fn)
    with open(os.path.join(d, "meta.txt"), "w") as f:
        for fn, r in zip(fns, rd):
            f.write(f"{fn}\t{r}\n")
    rec = np.zeros(len(inv), np.dtype([("i", "<u4"), ("c", "u1")]))
    rec["i"], rec["c"] = inv, ikmc
    with open(os.path.join(d, "ikmer.meta"), "wb") as f:
        f.write(np.array([nk, len(inv), n_tr], np.uint64).tobytes() + nk_cum.tobytes() + nik_cum.tobytes() + rec.tobytes())
    # ---- dbt_pred through the C ABI ----
    inp = abi.PredInput()
    inp.ns, inp.nk, inp.n_tr, inp.n_ik = ns, nk, n_tr, len(inv)
    arr = (abi.u64p * ns)(*[abi.ptr(v, abi.u64p) for v in kmc])
    inp.kmc = C.cast(arr, C.POINTER(abi.u64p))
    fp = C.POINTER(C.c_float)
    rdf = np.array([np.float32(float(f"{r}")) for r in rd], np.float32)
    inp.rd = rdf.ctypes.data_as(fp)
    inp.nk_cum, inp.nik_cum, inp.iki, inp.ikmc = abi.ptr(nk_cum, abi.u32p), abi.ptr(nik_cum, abi.u32p), abi.ptr(inv, abi.u32p), abi.ptr(ikmc, abi.u8p)
    raw, cor, bias = np.zeros((nk, ns), np.float32), np.zeros((nk, ns), np.float32), np.zeros((n_tr, ns), np.float32)
    ms = np.zeros(2, np.float32)
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        abi.check(L, L.dbt_pred(device, C.byref(inp), raw.ctypes.data_as(fp), cor.ctypes.data_as(fp), bias.ctypes.data_as(fp), ms.ctypes.data_as(fp)), "dbt_pred")
        wall = time.perf_counter() - t0
        if best is None or ms.sum() < best[0].sum():
            best = (ms.copy(), wall)
    ms, wall = [float(x) for x in best[0]], best[1]
    peak, peak_src = peaks()
    b_norm = ns * nk * 12
    b_bias = ns * nk * 8 + ns * len(inv) * 4 + ns * n_tr * 8
    blk = {"samples": ns, "kmers": nk, "loci": n_tr, "invariant_kmers": int(len(inv)), "dtype": "f32",
           "kernels": {"pred_norm_kernel": {"ms": float(ms[0]), "algorithmic_bytes": int(b_norm), "GBps": b_norm / ms[0] / 1e6, "frac_of_hbm_peak": b_norm / ms[0] / 1e6 / peak},
                       "pred_bias_kernel": {"ms": float(ms[1]), "algorithmic_bytes": int(b_bias), "GBps": b_bias / ms[1] / 1e6, "frac_of_hbm_peak": b_bias / ms[1] / 1e6 / peak}},
           "peak_GBps": peak, "peak_source": peak_src,
           "c_abi_call_s": wall, "c_abi_note": "dbt_pred with pageable host arrays: H2D of the counts, both kernels, D2H of the three matrices"}
    # ---- the two binaries on the same files ----
    def run(exe, tag):
        outs = [os.path.join(d, f"{tag}.{x}") for x in ("raw", "cor", "bias.tsv")]
        t0 = time.perf_counter()
        r = subprocess.run([exe, os.path.join(d, "meta.txt"), os.path.join(d, "ikmer.meta")] + outs, capture_output=True)
        if r.returncode:
            raise RuntimeError(f"{exe} failed: {r.stderr.decode()[-400:]}")
        ph = [l for l in r.stderr.decode().splitlines() if "phases:" in l]
        if ph:
            blk["cli_phases"] = ph[-1].split("phases:", 1)[1].strip()
        return time.perf_counter() - t0, outs
    t_gpu, o_gpu = run(os.path.join(ROOT, "danbing-tk_b200", "bin", "danbing-tk-pred"), "gpu")
    blk["cli_wall_s"] = t_gpu
    ref = os.path.join(ROOT, "oracle", "_ref", "danbing-tk-pred")
    if os.path.exists(ref):
        t_ref, o_ref = run(ref, "ref")
        same = [open(x, "rb").read() == open(y, "rb").read() for x, y in zip(o_gpu, o_ref)]
        blk["reference_binary"] = {"wall_s": t_ref, "threads": 1, "outputs_identical": {"raw": same[0], "corrected": same[1], "bias_tsv": same[2]}}
        blk["cli_speedup_vs_reference_binary"] = t_ref / t_gpu
        blk["abi_matches_cli"] = bool(np.array_equal(np.fromfile(o_gpu[1], np.float32, offset=8), cor.ravel()))
    shutil.rmtree(d, ignore_errors=True)
    return blk


def e2e_file_block(a, pref, pool, hap_off, workdir, device):
    """File to file (SURVEY §8(d) "end-to-end": FASTA bytes in the page cache -> outputs on disk): the drop-in CLI
    (danbing-tk_b200/bin/danbing-tk: raw text -> GPU record split + pairing -> pair loop -> kam / aln lines) on a FASTA of
    --file-pairs pairs in /dev/shm, and the unmodified reference binary on the same file with the same flags and all host cores.
    Query phase of both = wall time minus a zero-read run (index load); the CLI's own `--timing` query wall is reported beside it."""
    cli = os.path.join(ROOT, "danbing-tk_b200", "bin", "danbing-tk")
    n = a.file_pairs
    reads = np.empty(n * 2 * RLEN, np.uint8)
    synth_reads(device, pool, hap_off, 0, n, 4242, a.frac_locus, reads, SUB_PPM, INDEL_PPM)
    fa = os.path.join(workdir, "file.fa")
    write_fasta_sample(fa, reads, n)
    del reads
    empty = os.path.join(workdir, "file.empty.fa")
    open(empty, "w").close()
    nproc = os.cpu_count() or 1
    blk = {"pairs": n, "fasta_bytes": os.path.getsize(fa), "host_cores": nproc,
           "what": "FASTA in /dev/shm -> CLI -> stdout file + OUT.trkmc.ar + OUT.tr.summary.txt in /dev/shm; query phase = wall minus a zero-read run"}

    def timed(cmd, out_fn):
        t0 = time.perf_counter()
        with open(out_fn, "wb") as so:
            r = subprocess.run(cmd, stdout=so, stderr=subprocess.PIPE)
        if r.returncode:
            raise RuntimeError(f"{cmd[0]} failed: {r.stderr.decode()[-500:]}")
        return time.perf_counter() - t0, r.stderr.decode()

    def cli_run(flags, tag):
        base = [cli] + flags + ["-p", str(nproc), "--timing", "-qs", pref]
        t_load, _ = timed(base + ["-fa", empty, "-o", os.path.join(workdir, tag + ".e")], os.path.join(workdir, tag + ".e.out"))
        t_all, err = timed(base + ["-fa", fa, "-o", os.path.join(workdir, tag)], os.path.join(workdir, tag + ".out"))
        m = re.search(r"query wall ([0-9.]+) s; reader ([0-9.]+) s; upload\+split \(sync\) ([0-9.]+) s; pairer ([0-9.]+) s; wait ([0-9.]+) s; format ([0-9.]+) s; chunks (\d+) \((\d+) paired on the host\); launch ([0-9.]+) s; pairer turn wait ([0-9.]+) s; worker idle ([0-9.]+) s; reader blocked ([0-9.]+) s; stdout write ([0-9.]+) s", err)
        q = float(m.group(1)) if m else None
        d = {"pairs_per_s": n / max(t_all - t_load, 1e-9), "wall_s": t_all, "load_s": t_load, "query_wall_s_cli": q,
             "pairs_per_s_cli_query_wall": (n / q) if q else None, "stdout_bytes": os.path.getsize(os.path.join(workdir, tag + ".out"))}
        if m:
            d["phases_s"] = dict(zip(("reader", "upload+split", "pairer", "wait", "format"), map(float, m.groups()[1:6])))
            d["chunks"], d["chunks_paired_on_host"] = int(m.group(7)), int(m.group(8))
            d["phases_s"].update(dict(zip(("launch", "pairer_turn_wait", "worker_idle", "reader_blocked", "stdout_write"), map(float, m.groups()[8:13]))))
        return d
    asgn = ["-k", str(K), "-kf", "4", "1", "-cth", str(CTH)]
    blk["asgn"] = cli_run(asgn, "f_asgn")
    blk["thread"] = cli_run(asgn + (["-gc", "85", "3"] if CORR else ["-g", "85"]) + ["-ae"], "f_thr")
    blk["asgn_host_reader"] = cli_run(asgn + ["--host-reader"], "f_asgn_hr")
    exe = ref_bin("danbing-tk")
    if exe and not a.no_cpu:
        base = [exe] + asgn + ["-p", str(nproc), "-qs", pref]
        t_load, _ = timed(base + ["-fa", empty, "-o", os.path.join(workdir, "f_ref.e")], os.path.join(workdir, "f_ref.e.out"))
        t_all, _ = timed(base + ["-fa", fa, "-o", os.path.join(workdir, "f_ref")], os.path.join(workdir, "f_ref.out"))
        ref_rate = n / max(t_all - t_load, 1e-9)
        same = all(open(os.path.join(workdir, "f_asgn" + e), "rb").read() == open(os.path.join(workdir, "f_ref" + e), "rb").read()
                   for e in (".trkmc.ar", ".tr.summary.txt"))
        # stdout: the reference's threads print their kam lines in completion order; compare as sorted lines
        same_kam = sorted(open(os.path.join(workdir, "f_asgn.out"), "rb").read().split(b"\n")) == sorted(open(os.path.join(workdir, "f_ref.out"), "rb").read().split(b"\n"))
        blk["reference_binary"] = {"pairs_per_s": ref_rate, "wall_s": t_all, "load_s": t_load, "threads": nproc,
                                   "flags": "same as `asgn` (graph threading is dead code in the unmodified binary, SURVEY F2)",
                                   "outputs_identical": {"trkmc_ar_and_summary": same, "kam_lines_as_multiset": same_kam}}
        blk["speedup_vs_reference_binary"] = (blk["asgn"]["pairs_per_s_cli_query_wall"] or blk["asgn"]["pairs_per_s"]) / ref_rate
        blk["speedup_note"] = "CLI query wall (its own --timing line: first chunk read to last line written) vs the reference's wall minus its zero-read run"
    return blk


def hbm_resident_block(a, workdir):
    """The regime the "HBM random-access roofline" is about: same flags on an RPGG whose index and graph tables
    are far larger than the 126 MB L2 (--hbm-loci 20000: ~2 GB + ~2 GB), kernel-only, with a CPU-arm parity check on a sample."""
    import copy
    a2 = copy.copy(a)
    a2.loci, a2.pairs, a2.cpu_sample, a2.skip_e2e = a.hbm_loci, a.hbm_pairs, min(a.hbm_cpu_sample, a.hbm_pairs), True
    wd = os.path.join(workdir, "hbm")
    os.makedirs(wd, exist_ok=True)
    ix, pref, pool, hap_off = build_workload(a2.loci, 20261104, wd)
    r = b200_mode("thread", a2, ix, pref, pool, hap_off, wd, 0, 1, 0, None)
    rf = r["roofline"]
    blk = {"config": r["config"], "value": r["value"], "unit": r["unit"], "ms_per_step": r["ms_per_step"],
           "per_kernel_ms": rf["per_kernel_ms"], "r_rand": rf.get("r_rand"), "roofline_dominant": {k: rf[k] for k in ("kernel", "kernel_ms", "achieved", "frac")},
           "index_table": r["index_table"], "graph_table_bytes": r.get("graph_table_bytes"),
           "cpu_baseline": r["cpu_baseline"], "counters_per_step": r["counters_per_step_rank0"], "clocks": r["clocks"]}
    shutil.rmtree(wd, ignore_errors=True)
    return blk


def make_config(a, mode):
    err = f", locus reads with {SUB_PPM/1e4:.1f}% substitutions + {INDEL_PPM/1e4:.2f}% indels" if mode == "thread" else ", error-free"
    return {"workload": f"synthetic {a.loci}-locus RPGG (8 haplotypes, 700 bp flanks), {a.pairs} simulated 150bp read pairs per GPU "
                        f"({a.frac_locus:.0%} locus reads{err} / rest background), {flags_txt(mode)}",
            "mode": mode, "pairs_per_gpu": a.pairs, "loci": a.loci, "k": K, "cth": CTH, "kf": [4, 1],
            "l2_policy": "inputs larger than L2 (packed reads 1.3 GB/step >> 126 MB); no explicit flush"}


def reference_arm(a, workdir):
    mode = "asgn" if a.mode == "asgn" else "thread"
    exe = ref_bin("ref_harness" if mode == "thread" else "danbing-tk")
    if not exe:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"}))
        return
    # nothing of the product is loaded here (no libdanbing_b200.so, no CUDA): the reads come from the host build of the
    # workload generator and are byte-identical to what rank 0 of the B200 arm generates on its GPU
    nproc = os.cpu_count() or 1
    ix, pref, pool, hap_off = build_workload(a.loci, 20261104, workdir)
    n = min(a.cpu_sample, a.pairs)
    reads = np.empty(n * 2 * RLEN, np.uint8)
    e = (SUB_PPM, INDEL_PPM) if mode == "thread" else (0, 0)
    synth_reads_host(pool, hap_off, 0, n, 777, a.frac_locus, reads, *e)
    times = []
    if mode == "asgn":
        fa = os.path.join(workdir, "sample.fa")
        write_fasta_sample(fa, reads, n)
    used = nproc
    for i in range(a.warmup + a.steps):
        if mode == "thread":
            t_q, t_all, t_load, _, used = run_reference_thread(pref, reads, n, workdir, nproc)
        else:
            t_q, t_all, t_load = run_reference(pref, fa, os.path.join(workdir, "ref"), nproc)
        if i >= a.warmup:
            times.append(t_q)
    ms = 1e3 * sum(times) / len(times)
    v = n / (ms / 1e3)
    how = ("reference functions (isThreadFeasible & helpers) compiled from /root/reference as oracle/_ref/ref_harness, caller cpp:2072-2088 restored (D1), "
           f"{used} processes on shards of the sample (pairs as TSV in /dev/shm)") if mode == "thread" else \
          f"unmodified oracle/_ref/danbing-tk -p {nproc}, FASTA in /dev/shm"
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "read pairs/s", "n_gpus": a.gpus,
                      "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": make_config(a, mode),
                      "cpu_baseline": {"value": v, "unit": "read pairs/s", "cores": used, "kind": "reference",
                                       "sample": f"first {n} pairs of the workload; {how}; query phase = wall minus a zero-read (index load) run"},
                      "e2e": {"value": v, "unit": "read pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def reference_tr_order(pref, workdir):
    """(tr_n, tr_kmers) of PREF.tr.kmers in the reference's OUTPUT order, via the product's host loader (no GPU code)."""
    exe = os.path.join(ROOT, "danbing-tk_b200", "bin", "dbt-hostcheck")
    out = os.path.join(workdir, "tr.order.u64")
    if not os.path.exists(exe) or subprocess.run([exe, "tr-order", pref + ".tr.kmers", out]).returncode:
        return None
    b = np.fromfile(out, "<u8")
    nloci = int(b[0])
    return b[1:1 + nloci].copy(), b[1 + nloci:].copy()


def per_locus_multiset_equal(ix, a, b):
    ro = np.concatenate([[0], np.cumsum(ix["tr_n"]).astype(np.int64)])
    for l in range(len(ix["tr_n"])):
        if not np.array_equal(np.sort(a[ro[l]:ro[l + 1]]), np.sort(b[ro[l]:ro[l + 1]])):
            return False
    return True


_numa_bound = False
_numa_info = {"bound_to_gpu_local_cpus": False}


def bind_to_gpu_numa_node(local):
    """Pin this rank's threads to the CPUs next to its GPU before any pinned host memory is allocated, so that the
    staging buffers land on the GPU's own NUMA node (matters at 8 ranks: 8 x 52 GB/s of H2D out of host DRAM)."""
    global _numa_bound
    if _numa_bound:
        return
    _numa_bound = True
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        cpus = sorted(os.sched_getaffinity(0))
        _numa_info.update({"bound_to_gpu_local_cpus": True, "ncpus": len(cpus), "cpus": f"{cpus[0]}-{cpus[-1]}" if cpus else ""})
        log(f"[bench] rank-local GPU {local}: bound to {len(cpus)} NUMA-local CPUs")
    except Exception as e:                                    # affinity is an optimisation only
        log(f"[bench] NUMA binding skipped: {e}")


def b200_mode(mode, a, ix, pref, pool, hap_off, workdir, rank, world, local, dist):
    thr = mode == "thread"
    if world > 1:
        bind_to_gpu_numa_node(local)
    L = abi.load_library()
    if L.dbt_device_count() <= local:
        raise RuntimeError("bench.py needs a CUDA device per rank: libdanbing_b200 has no CPU fallback")
    config = make_config(a, mode)
    P = a.pairs
    flags = (abi.F_THREADING | abi.F_CORRECTION | abi.F_ALN | abi.F_ALN_MIN) if thr else abi.F_KAM
    if thr and not CORR:
        flags &= ~abi.F_CORRECTION
    p = abi.default_params(k=K, cth=CTH, n_filter=4, nm_filter=1, nm_tr=40, flags=flags, max_read_len=RLEN,
                           max_batch_pairs=P, device=local)
    p.thread_cth, p.max_ncorrection = 85, 3
    p.max_records = max(int(P * a.frac_locus * 1.5) + 4096, 65536)
    p.bucket_slots = a.bucket_slots
    p.table_load = a.table_load
    t = time.time()
    ixm = ix if thr else {k: v for k, v in ix.items() if not k.startswith("g_")}
    ctx = abi.Context(p, abi.make_index_struct(ixm))
    if thr and not ctx.threading:
        raise RuntimeError("threading mode requested but the context did not enable restored threading")
    log(f"[bench r{rank}] {mode}: index resident in HBM ({time.time()-t:.1f}s)")
    stats = np.zeros(8, np.uint64)
    abi.check(L, L.dbt_table_stats(ctx.h, abi.ptr(stats, abi.u64p)), "table_stats")
    if world > 1:
        import torch
        idt = torch.zeros(abi.DBT_NCCL_ID_BYTES, dtype=torch.uint8)
        if rank == 0:
            idb = np.zeros(abi.DBT_NCCL_ID_BYTES, np.uint8)
            abi.check(L, L.dbt_nccl_unique_id(abi.ptr(idb, abi.u8p)), "nccl id")
            idt = torch.from_numpy(idb)
        dist.broadcast(idt, 0)
        idb = idt.numpy().copy()
        sys.stdout.flush()
        keep = os.dup(1)                                       # NCCL prints its version banner on fd 1: keep stdout to the one JSON line
        os.dup2(2, 1)
        try:
            abi.check(L, L.dbt_comm_init(ctx.h, world, rank, abi.ptr(idb, abi.u8p)), "comm init")
            abi.check(L, L.dbt_allreduce(ctx.h), "allreduce")    # first collective (lazy channel set-up) outside the timed region; synchronous
        finally:
            os.dup2(keep, 1)
            os.close(keep)

    # ---- synthetic reads of this rank: pinned host ASCII (for e2e) ----
    nbytes = P * 2 * RLEN
    host, owner = abi.pinned_array(L, nbytes)
    t = time.time()
    synth_reads(local, pool, hap_off, rank * P, P, 777, a.frac_locus, host, *((SUB_PPM, INDEL_PPM) if thr else (0, 0)))
    offsets_np = np.arange(2 * P + 1, dtype=np.uint64) * RLEN
    off_pin, off_owner = abi.pinned_array(L, offsets_np.nbytes)
    off_pin.view(np.uint64)[:] = offsets_np
    offsets = off_pin.view(np.uint64)
    log(f"[bench r{rank}] {mode}: {P} pairs generated ({time.time()-t:.1f}s)")

    def barrier():
        if dist is not None:
            dist.barrier()

    def allred(x, op):
        if dist is None:
            return x
        import torch
        tt = torch.tensor([float(x)], dtype=torch.float64)
        dist.all_reduce(tt, op=op)
        return float(tt[0])

    def allmax(x):
        return allred(x, dist.ReduceOp.MAX) if dist is not None else x

    def allsum(x):
        return allred(x, dist.ReduceOp.SUM) if dist is not None else x

    # ---- 1. kernel-only: reads resident in HBM ----
    batch = ctx.make_batch(host, offsets)
    abi.check(L, L.dbt_upload(ctx.h, C.byref(batch), 0), "upload")
    abi.check(L, L.dbt_wait(ctx.h, 0), "wait")
    res = abi.BatchResult()
    # pack_kernel (ASCII -> 2-bit records) of this upload: 2P reads, CUDA events around the launch
    kt0 = np.zeros(8, np.float32)
    kl0 = np.zeros(8, np.uint32)
    abi.check(L, L.dbt_kernel_times(ctx.h, 0, kt0.ctypes.data_as(C.POINTER(C.c_float)), abi.ptr(kl0, abi.u32p)), "ktimes")
    t_pack = float(kt0[0])
    bytes_pack = 2 * P * (RLEN + 8 + int(L.dbt_packed_bytes(RLEN, 1)) // 2)      # ASCII + u64 offset in, packed record + meta out

    def step():
        abi.check(L, L.dbt_process(ctx.h, 0), "process")
        if world > 1:
            abi.check(L, L.dbt_allreduce(ctx.h), "allreduce")

    for _ in range(a.warmup):
        step()
    abi.check(L, L.dbt_wait(ctx.h, 0), "wait")
    launches0 = L.dbt_launch_count(ctx.h)
    ktimes = np.zeros((a.steps, 8), np.float32)
    kl = np.zeros(8, np.uint32)
    barrier()
    with ClockSampler(local) as clk:
        ms = C.c_float()
        abi.check(L, L.dbt_timer_start(ctx.h, 0), "timer")
        t0 = time.perf_counter()
        for i in range(a.steps):
            step()
            if a.steps <= 64:        # per-kernel CUDA-event times of this step (synchronises the slot)
                abi.check(L, L.dbt_kernel_times(ctx.h, 0, ktimes[i].ctypes.data_as(C.POINTER(C.c_float)), abi.ptr(kl, abi.u32p)), "ktimes")
        abi.check(L, L.dbt_timer_stop(ctx.h, 0, C.byref(ms)), "timer")
        wall_ms = (time.perf_counter() - t0) * 1e3
        barrier()
        time.sleep(0.3)
    launches = int(L.dbt_launch_count(ctx.h) - launches0)
    # CUDA events on the slot's stream at every N: dbt_allreduce is host-synchronous, so the next step's kernels are enqueued
    # after it returned and the reduce shows up as a gap between the two events
    dev_ms = float(ms.value)
    step_ms = allmax(dev_ms) / a.steps
    wall_step_ms = allmax(wall_ms) / a.steps
    value = world * P / (step_ms / 1e3)
    # counters of one step (fetch once, untimed)
    abi.check(L, L.dbt_fetch(ctx.h, C.byref(res), 0), "fetch")
    abi.check(L, L.dbt_wait(ctx.h, 0), "wait")
    cnt = dict(zip(abi.COUNTER_NAMES, list(res.counters)))

    # ---- multi-GPU parity (untimed): ONE NCCL reduce of one step's results == the sum of the ranks' own arrays ----
    # reset -> one step per rank -> download the rank-local arrays -> dbt_allreduce once -> download the reduced arrays;
    # the expected value is a gloo all_reduce(sum) of the local arrays (u64 arithmetic wraps the same way in int64)
    mgp = None
    if world > 1:
        import torch
        ctx.reset()
        abi.check(L, L.dbt_process(ctx.h, 0), "process")
        abi.check(L, L.dbt_fetch(ctx.h, C.byref(res), 0), "fetch")
        abi.check(L, L.dbt_wait(ctx.h, 0), "wait")
        loc = ctx.download()
        abi.check(L, L.dbt_allreduce(ctx.h), "allreduce")
        red = ctx.download()
        names = ("trkmc", "kmc", "nmapread", "counters")
        ok = {}
        for nm, lo_, re_ in zip(names, loc, red):
            if nm == "nmapread":
                t = torch.from_numpy(lo_.astype(np.int64))
                dist.all_reduce(t)
                want = (t.numpy() & 0xFFFFFFFF).astype(np.uint32)         # atomic_uint32_t in the reference (cpp:1778): wraps
            else:
                t = torch.from_numpy(np.ascontiguousarray(lo_, np.uint64).view(np.int64).copy())
                dist.all_reduce(t)
                want = t.numpy().view(np.uint64)
            ok[nm] = bool(np.array_equal(want, re_))
        flag = torch.tensor([int(all(ok.values()))])
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        tot_red = int(red[0].sum())
        mgp = {"nccl_equals_sum_of_ranks": bool(int(flag[0])), "per_array_rank0": ok, "nranks": world,
               "reduced_trkmc_total": tot_red, "rank0_local_trkmc_total": int(loc[0].sum()),
               "how": "reset, one step per rank, dbt_allreduce once (out of place), compared with a gloo all_reduce(sum) of the ranks' "
                      "pre-reduce trkmc / kmc / nmapread / counters; MIN over ranks of the per-rank verdict"}
        ctx.reset()
    kt = ktimes.mean(axis=0)
    t_sub, t_cmp, t_cand, t_gen, t_thr, t_thr2 = float(kt[1]), float(kt[2]), float(kt[3]), float(kt[4]), float(kt[5]), float(kt[6])

    # ---- roofline of the dominant kernel (algorithmic bytes: SURVEY.md §8(d)) ----
    # one logical probe = one 32 B sector, one count increment = 64 B read-modify-write, bases at 2 bits
    nk = RLEN - K + 1
    ncand = cnt["ncandidates"]
    p_sub = cnt["nhash0"] + 2 * ncand                       # ++nhash skips the iteration that breaks (cpp:176)
    bytes_sub = P * 75 + 32 * p_sub
    g_h1, g_thr, g_inc = cnt["gen_nhash1"], cnt["gen_nThreadingReads"], cnt["gen_ninc"]
    merged = thr and L.dbt_graph_merged(ctx.h) == 1
    if thr:
        # candidate stage = kfilter + vote only (no assignTRkmc); threading: one node look-up per k-mer step of the walk
        # (cpp:1170), the look-ups of the correction code, one trResults look-up per k-mer of the edited vectors of
        # counted pairs (cpp:2203-2208) and the increments.  All four are the reference algorithm's own counts.
        bytes_gen = 32 * g_h1
        bytes_fast = 32 * (cnt["nhash1"] - g_h1)
        # split between the two threading kernels by thread_kernel's own counters (thr2_*): thread_fast_kernel walks and
        # classifies every error-free mate, thread_kernel the mates with errors (+ the correction look-ups)
        bytes_thr2 = 32 * (cnt["thr2_walk"] + cnt["thr_ec"] + cnt["thr2_cls"]) + 64 * cnt["thr2_ninc"]
        bytes_thr = 32 * (cnt["thr_walk"] + cnt["thr_ec"] + cnt["thr_cls"]) + 64 * cnt["ninc"] - bytes_thr2
        if merged:
            # graphs merged into the index entries: candidate_fast_kernel does thread_fast_kernel's work itself (walk + classification of
            # every mate it can finish, their increments) out of the index buckets it has probed; thread_fast_kernel only sees the pairs
            # the generic kernel defers (0.3 % of the candidates; its own counters are not kept apart, its share is booked with the rest)
            bytes_fast += bytes_thr
            bytes_thr = 0
    else:
        # candidate stage: kfilter probes (nhash1) + one classification probe per k-mer of pairs reaching assignTRkmc + increments,
        # split between the closed-form kernel and the generic kernel by the generic kernel's own counters
        bytes_gen = 32 * (g_h1 + g_thr * nk) + 64 * g_inc
        bytes_fast = 32 * ((cnt["nhash1"] - g_h1) + (cnt["nThreadingReads"] - g_thr) * nk) + 64 * (cnt["ninc"] - g_inc)
        bytes_thr = bytes_thr2 = 0
    peak, peak_src = peaks()
    kern = {"subfilter_kernel": (t_sub, bytes_sub), "candidate_fast_kernel": (t_cand, bytes_fast), "candidate_kernel(generic)": (t_gen, bytes_gen)}
    if thr:
        kern["thread_fast_kernel"] = (t_thr, bytes_thr)
        kern["thread_kernel"] = (t_thr2, bytes_thr2)
    dom = max(kern, key=lambda n: kern[n][0])
    dom_ms, dom_bytes = kern[dom]
    ach = dom_bytes / (dom_ms / 1e3) / 1e9 if dom_ms > 0 else 0.0
    step_bytes = bytes_sub + bytes_fast + bytes_gen + bytes_thr + bytes_thr2
    roofline = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                "kernel": dom, "kernel_ms": dom_ms, "algorithmic_bytes_per_launch": dom_bytes, "peak_source": peak_src,
                "per_kernel_ms": {"subfilter": t_sub, "compaction": t_cmp, "candidate_fast": t_cand, "candidate_generic": t_gen, "thread_fast": t_thr, "thread": t_thr2,
                                  "pack(e2e upload only, 2P reads)": t_pack},
                "per_kernel_GBps": dict({n: (bb / (ms_ / 1e3) / 1e9 if ms_ > 0 else 0.0) for n, (ms_, bb) in kern.items()},
                                        pack_kernel=(bytes_pack / (t_pack / 1e3) / 1e9 if t_pack > 0 else 0.0)),
                "pack_kernel": {"ms": t_pack, "reads": 2 * P, "bytes": bytes_pack, "GBps": bytes_pack / (t_pack / 1e3) / 1e9 if t_pack > 0 else 0.0,
                                "frac_of_stream_peak": (bytes_pack / (t_pack / 1e3) / 1e9 / peak) if t_pack > 0 else 0.0,
                                "note": "ASCII -> 2-bit records; runs once per ASCII upload (e2e path), not inside the kernel-only step"},
                "per_kernel_bytes": {n: bb for n, (ms_, bb) in kern.items()},
                "generic_path_pairs": cnt["gen_pairs"], "candidate_pairs": ncand,
                "step_bytes": step_bytes, "step_GBps": step_bytes / (step_ms / 1e3) / 1e9, "step_frac": step_bytes / (step_ms / 1e3) / 1e9 / peak}
    if thr:
        nthr_pairs = max(cnt["nThreadingReads"] // 2, 1)
        t_stage = t_thr + t_thr2
        roofline["graphs_merged_into_index"] = bool(merged)
        roofline["thread_stage"] = {"kernels": "thread_fast_kernel + thread_kernel (= round 1's single thread_kernel)" if not merged else
                                    "thread_kernel only: with the graphs merged into the index candidate_fast_kernel threads the error-free / corrected mates itself "
                                    "(its time and bytes are booked under candidate_fast_kernel); thread_fast_kernel takes the pairs the generic kernel defers", "ms": t_stage,
                                    "algorithmic_bytes": bytes_thr + bytes_thr2,
                                    "GBps": (bytes_thr + bytes_thr2) / (t_stage / 1e3) / 1e9 if t_stage > 0 else 0.0,
                                    "frac": (bytes_thr + bytes_thr2) / (t_stage / 1e3) / 1e9 / peak if t_stage > 0 else 0.0,
                                    "pairs": cnt["nThreadingReads"] // 2, "pairs_with_an_error_mate": cnt["thr2_pairs"], "mates_with_errors": cnt["thr2_mates"],
                                    "probes_algorithmic_per_pair": (cnt["thr_walk"] + cnt["thr_ec"] + cnt["thr_cls"]) / nthr_pairs,
                                    "probes_issued_per_pair": (cnt["thr_refresh"] + cnt["thr_ec"]) / nthr_pairs,
                                    "Gprobes_per_s": (cnt["thr_refresh"] + cnt["thr_ec"]) / (t_stage / 1e3) / 1e9 if t_stage > 0 else 0.0}
    # dram bytes per launch of the dominant kernel from the committed `ncu --set full` capture of this same workload
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        w = tj.get(mode, tj).get("workload", {})
        tm = tj.get(mode, tj)
        if (w.get("pairs"), w.get("loci"), w.get("frac_locus")) == (P, a.loci, a.frac_locus) and dom in tm:
            roofline["traffic"] = tm[dom]
            roofline["traffic_source"] = "profiles/traffic.json (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum)"
    except (OSError, ValueError):
        pass
    if a.rrand and rank == 0:
        rr = {}
        # R_rand depends on HOW a sector is asked for (profiles/r2/k4_rrand_sweep.json): two 16-byte loads of one thread, one 256-bit
        # load per thread (what the probe functions issue), or one coalesced request from adjacent lanes.  The denominator of every
        # fraction below is the BEST of them -- the rate the memory system can deliver, not the rate of one access pattern.
        def gather_rate(mode_bits, sect):
            msr = C.c_float()
            abi.check(L, L.dbt_random_gather_bench(local, 8 << 30, sect | mode_bits, 1 << 28, 3, C.byref(msr)), "rrand")
            return (1 << 28) / (msr.value / 1e3) / 1e9
        rr["32B_by_access_pattern"] = {"two_16B_loads_per_thread": gather_rate(0, 32), "one_256bit_load_per_thread": gather_rate(0x20000, 32),
                                       "one_coalesced_request_from_2_lanes": gather_rate(0x10000, 32)}
        rr["128B_line_one_coalesced_request_Glines_per_s"] = gather_rate(0x10000, 128)
        rr["32B_Gsectors_per_s"] = max(rr["32B_by_access_pattern"].values())
        rr["64B_Gsectors_per_s"] = max(gather_rate(0, 64), gather_rate(0x10000, 64))
        # the "HBM random-access roofline" of SURVEY.md 8(d): logical probes (one 32 B sector each) per second of the dominant
        # kernel against the measured gather rate over a table far beyond L2.  At the default index size most probes hit L2, so
        # this can exceed 1 -- it is the honest denominator for HBM-resident tables (--loci 20000 and up).
        # logical = the probes the REFERENCE algorithm makes (SURVEY §8(d) unit of work); issued = the bucket fetches the kernel
        # really starts (one graph probe answers walk + classification; kfilter and assignTRkmc share one index probe)
        l_tf = (cnt["thr_walk"] - cnt["thr2_walk"]) + (cnt["thr_cls"] - cnt["thr2_cls"]) if thr else 0
        i_tf = cnt["thr_refresh"] - cnt["thr2_refresh"] if thr else 0
        logical = {"thread_fast_kernel": 0 if merged else l_tf,
                   "thread_kernel": cnt["thr2_walk"] + cnt["thr_ec"] + cnt["thr2_cls"] if thr else 0,
                   "candidate_fast_kernel": (cnt["nhash1"] - g_h1) + (0 if thr else (cnt["nThreadingReads"] - g_thr) * nk) + (l_tf if merged else 0),
                   "subfilter_kernel": p_sub}
        issued = {"thread_fast_kernel": 0 if merged else i_tf,
                  "thread_kernel": cnt["thr2_refresh"] + cnt["thr_ec"] if thr else 0,
                  "candidate_fast_kernel": cnt["nhash1"] - g_h1 + (i_tf if merged else 0),
                  "subfilter_kernel": p_sub}
        if logical.get(dom) and dom_ms > 0:
            rr["dominant_kernel_Gprobes_per_s"] = logical[dom] / (dom_ms / 1e3) / 1e9
            rr["frac_of_r_rand_32B"] = rr["dominant_kernel_Gprobes_per_s"] / rr["32B_Gsectors_per_s"]
            rr["dominant_kernel_issued_Gprobes_per_s"] = issued[dom] / (dom_ms / 1e3) / 1e9
            rr["frac_of_r_rand_32B_issued"] = rr["dominant_kernel_issued_Gprobes_per_s"] / rr["32B_Gsectors_per_s"]
        tk = {"thread_fast_kernel": t_thr, "thread_kernel": t_thr2, "candidate_fast_kernel": t_cand, "subfilter_kernel": t_sub}
        rr["per_kernel"] = {n: {"logical_Gprobes_per_s": logical[n] / (tk[n] / 1e3) / 1e9, "issued_Gprobes_per_s": issued[n] / (tk[n] / 1e3) / 1e9,
                                "frac_logical": logical[n] / (tk[n] / 1e3) / 1e9 / rr["32B_Gsectors_per_s"],
                                "frac_issued": issued[n] / (tk[n] / 1e3) / 1e9 / rr["32B_Gsectors_per_s"]}
                            for n in tk if tk[n] > 0 and logical[n]}
        # T_min = bytes_streamed / BW_stream + sectors_random / R_rand; fraction = T_min / T_measured (the whole step)
        streamed = P * 75 + (cnt["nThreadingReads"] // 2) * (ctx.stride if thr else 0)
        t_meas = step_ms / 1e3
        for nm, sect in (("issued", sum(issued.values())), ("logical", sum(logical.values()))):
            sect += 2 * cnt["ninc"]                               # a count increment = 64 B read-modify-write = two sectors
            t_min = streamed / (peak * 1e9) + sect / (rr["32B_Gsectors_per_s"] * 1e9)
            rr[f"step_T_min_ms_{nm}"] = t_min * 1e3
            rr[f"step_frac_{nm}"] = t_min / t_meas if t_meas > 0 else 0.0
        rr["step_T_measured_ms"] = step_ms
        roofline["r_rand"] = rr

    e2e = None
    if not getattr(a, "skip_e2e", False):
        # ---- 2. end to end through the C ABI with host buffers (double-buffered sub-batches) ----
        sb = min(a.sub_batch, P)
        nsub = (P + sb - 1) // sb
        stride = ctx.stride
        cap_sb = min(p.max_records, int(sb * a.frac_locus * 1.5) + 4096)
        recpin = [abi.pinned_array(L, cap_sb * stride) for _ in range(2)]      # pinned result buffers (dbt_host_alloc): records arrive by DMA
        recbuf = [x[0] for x in recpin]
        results = [abi.BatchResult() for _ in range(2)]
        d2h = [0]

        packed_sb = L.dbt_packed_bytes(RLEN, sb)
        packpin = [abi.pinned_array(L, packed_sb) for _ in range(2)] if a.e2e_input in ("hostpack", "both") else None
        # hybrid: PCIe carries ASCII for most sub-batches while the host cores pack every `hybrid_every`-th one (64 B/read on the bus)
        sbh = max(1, sb // 2)
        nsubh = (P + sbh - 1) // sbh
        hyb = [j for j in range(nsubh) if j % a.hybrid_every == a.hybrid_every - 1] if a.e2e_input in ("hybrid", "both") else []
        hybpin = {j: abi.pinned_array(L, L.dbt_packed_bytes(RLEN, sbh)) for j in hyb}
        pack_threads = a.pack_threads if a.pack_threads > 0 else len(os.sched_getaffinity(0))
        pack_s = [0.0]

        def e2e_step(hostpack):
            pending = [False, False]
            nrec = 0
            for j in range(nsub):
                s = j & 1
                if pending[s]:
                    abi.check(L, L.dbt_wait(ctx.h, s), "wait")
                    nrec += int(results[s].n_records)
                b0, b1 = j * sb, min(P, (j + 1) * sb)
                if hostpack:
                    # host side of the batcher: ASCII -> 2-bit records in a pinned staging buffer (all host threads, AVX2), while the
                    # GPU works on the previous sub-batch; the H2D copy then moves 64 B per read instead of 150 B + 8 B of offset
                    tp = time.perf_counter()
                    abi.check(L, L.dbt_pack_reads(K, RLEN, b1 - b0, abi.ptr(host, abi.u8p), abi.ptr(offsets[2 * b0: 2 * b1 + 1], abi.u64p),
                                                  abi.ptr(packpin[s][0], abi.u8p), pack_threads), "pack")
                    pack_s[0] += time.perf_counter() - tp
                    bb = ctx.make_batch(packpin[s][0], None, first_pair_index=rank * P + b0, layout=abi.LAYOUT_PACKED, npairs=b1 - b0)
                else:
                    bb = ctx.make_batch(host, offsets[2 * b0: 2 * b1 + 1], first_pair_index=rank * P + b0)
                results[s].capacity = cap_sb
                results[s].records = abi.ptr(recbuf[s], abi.u8p)
                abi.check(L, L.dbt_submit(ctx.h, C.byref(bb), C.byref(results[s]), s), "submit")
                pending[s] = True
            for s in range(2):
                if pending[s]:
                    abi.check(L, L.dbt_wait(ctx.h, s), "wait")
                    nrec += int(results[s].n_records)
            if world > 1:
                abi.check(L, L.dbt_allreduce(ctx.h), "allreduce")
            d2h[0] = nrec * stride + nsub * (abi.DBT_N_COUNTERS * 8 + 8)
            return nrec

        def e2e_step_hybrid():
            import threading
            done = {j: threading.Event() for j in hyb}
            rcs = []

            def packer():
                for j in hyb:
                    b0, b1 = j * sbh, min(P, (j + 1) * sbh)
                    tp = time.perf_counter()
                    rcs.append(L.dbt_pack_reads(K, RLEN, b1 - b0, abi.ptr(host, abi.u8p), abi.ptr(offsets[2 * b0: 2 * b1 + 1], abi.u64p),
                                                abi.ptr(hybpin[j][0], abi.u8p), max(1, pack_threads - 1)))
                    pack_s[0] += time.perf_counter() - tp
                    done[j].set()

            th = threading.Thread(target=packer)
            th.start()
            pending = [False, False]
            nrec = 0
            for j in range(nsubh):
                s = j & 1
                if pending[s]:
                    abi.check(L, L.dbt_wait(ctx.h, s), "wait")
                    nrec += int(results[s].n_records)
                b0, b1 = j * sbh, min(P, (j + 1) * sbh)
                if j in done:
                    done[j].wait()
                    bb = ctx.make_batch(hybpin[j][0], None, first_pair_index=rank * P + b0, layout=abi.LAYOUT_PACKED, npairs=b1 - b0)
                else:
                    bb = ctx.make_batch(host, offsets[2 * b0: 2 * b1 + 1], first_pair_index=rank * P + b0)
                results[s].capacity = cap_sb
                results[s].records = abi.ptr(recbuf[s], abi.u8p)
                abi.check(L, L.dbt_submit(ctx.h, C.byref(bb), C.byref(results[s]), s), "submit")
                pending[s] = True
            for s in range(2):
                if pending[s]:
                    abi.check(L, L.dbt_wait(ctx.h, s), "wait")
                    nrec += int(results[s].n_records)
            th.join()
            if any(rcs):
                raise RuntimeError("dbt_pack_reads failed in the hybrid e2e path")
            if world > 1:
                abi.check(L, L.dbt_allreduce(ctx.h), "allreduce")
            d2h[0] = nrec * stride + nsubh * (abi.DBT_N_COUNTERS * 8 + 8)
            return nrec

        def e2e_run(hostpack):
            ctx.reset()
            step_fn = e2e_step_hybrid if hostpack == "hybrid" else (lambda: e2e_step(hostpack))
            for _ in range(min(a.warmup, 2)):
                step_fn()
            ctx.reset()
            pack_s[0] = 0.0
            barrier()
            t0 = time.perf_counter()
            for _ in range(a.steps):
                nrec = step_fn()
            e2e_ms = allmax((time.perf_counter() - t0) * 1e3) / a.steps
            if hostpack == "hybrid":
                hb = sum(L.dbt_packed_bytes(RLEN, min(P, (j + 1) * sbh) - j * sbh) for j in hyb)
                ab = sum((min(P, (j + 1) * sbh) - j * sbh) * (2 * RLEN + 16) for j in range(nsubh) if j not in hybpin) + 8 * (nsubh - len(hyb))
                return {"value": world * P / (e2e_ms / 1e3), "unit": "read pairs/s", "h2d_bytes_per_step": int(hb + ab),
                        "d2h_bytes_per_step": int(d2h[0]), "ms_per_step": e2e_ms, ("aln_records_per_step" if thr else "kam_records_per_step"): nrec,
                        "record_bytes": int(stride),
                        "input": ("pinned host ASCII reads + u64 offsets; sub-batches %s of %d packed to 2 bits by dbt_pack_reads on the host (%d threads, "
                                  "inside the timed region, overlapped with the ASCII H2D of the others), the rest packed on the device"
                                  % (hyb, nsubh, max(1, pack_threads - 1))),
                        "sub_batch_pairs": sbh, "host_pack_ms_per_step": pack_s[0] * 1e3 / a.steps}
            r = {"value": world * P / (e2e_ms / 1e3), "unit": "read pairs/s",
                 "h2d_bytes_per_step": int(nsub and sum(L.dbt_packed_bytes(RLEN, min(P, (j + 1) * sb) - j * sb) for j in range(nsub))) if hostpack
                 else int(nbytes + offsets_np.nbytes),
                 "d2h_bytes_per_step": int(d2h[0]), "ms_per_step": e2e_ms, ("aln_records_per_step" if thr else "kam_records_per_step"): nrec,
                 "record_bytes": int(stride),
                 "input": ("pinned host ASCII reads + u64 offsets -> dbt_pack_reads on the host (AVX2, %d threads, inside the timed region) -> "
                           "2-bit records H2D" % pack_threads) if hostpack else "pinned host ASCII reads + u64 offsets, packed to 2 bits on the device",
                 "sub_batch_pairs": sb}
            if hostpack:
                r["host_pack_ms_per_step"] = pack_s[0] * 1e3 / a.steps
                r["host_pack_GBps_ascii"] = nbytes / (pack_s[0] / a.steps) / 1e9 if pack_s[0] > 0 else None
            return r

        # both input paths start from the same pinned ASCII reads and end with the same records in host memory; the faster one on
        # this host is the e2e figure, the other is reported beside it
        runs = {}
        if a.e2e_input in ("ascii", "both"):
            runs["ascii"] = e2e_run(False)
        if a.e2e_input in ("hostpack", "both"):
            runs["hostpack"] = e2e_run(True)
        if a.e2e_input in ("hybrid", "both") and hyb:
            runs["hybrid"] = e2e_run("hybrid")
        best = max(runs, key=lambda n: runs[n]["value"])
        e2e = dict(runs[best])
        e2e["path"] = best
        # which resource the e2e step saturates: the bus (against this box's own pinned-copy rate, measured now), the host
        # packer cores (hybrid / hostpack: packer busy time against the step), or neither (the kernels)
        h2d_ms, d2h_ms = C.c_float(0), C.c_float(0)
        abi.check(L, L.dbt_bus_bench(local, 1 << 30, 3, C.byref(h2d_ms), C.byref(d2h_ms)), "bus bench")
        bus_h2d, bus_d2h = (1 << 30) / h2d_ms.value / 1e6, (1 << 30) / d2h_ms.value / 1e6
        ach_h2d, ach_d2h = e2e["h2d_bytes_per_step"] / e2e["ms_per_step"] / 1e6, e2e["d2h_bytes_per_step"] / e2e["ms_per_step"] / 1e6
        packer_busy = e2e.get("host_pack_ms_per_step", 0.0) / e2e["ms_per_step"]
        e2e["saturated"] = {"resource": ("PCIe host->device" if ach_h2d >= 0.8 * bus_h2d else "host packer cores" if packer_busy >= 0.8 else "GPU kernels / launch latency"),
                            "h2d_GBps": ach_h2d, "h2d_GBps_box_peak_pinned_1GiB_copy": bus_h2d, "h2d_frac": ach_h2d / bus_h2d,
                            "d2h_GBps": ach_d2h, "d2h_GBps_box_peak_pinned_1GiB_copy": bus_d2h,
                            "host_packer_busy_frac": packer_busy, "host_cores": len(os.sched_getaffinity(0)),
                            "gpu_step_ms_inputs_resident": step_ms}
        nkey = "aln_records_per_step" if thr else "kam_records_per_step"
        e2e["other_paths"] = []
        for n, r in runs.items():
            if n != best:
                e2e["other_paths"].append({"path": n, "value": r["value"], "ms_per_step": r["ms_per_step"], "h2d_bytes_per_step": r["h2d_bytes_per_step"]})
                if r.get(nkey) != e2e.get(nkey):
                    raise RuntimeError("e2e input paths disagree on the number of records")

    # ---- 3. CPU reference on a bounded sample of the same reads, outputs diffed against the GPU ----
    cpu = None
    if rank == 0 and not a.no_cpu:
        nproc = os.cpu_count() or 1
        n = min(a.cpu_sample, P)
        if ref_bin("ref_harness" if thr else "danbing-tk"):
            # GPU on the same sample, compared through the reference's own output files
            p2 = abi.default_params(k=K, cth=CTH, flags=(flags & ~(abi.F_ALN | abi.F_ALN_MIN | abi.F_KAM)), max_read_len=RLEN, max_batch_pairs=n, device=local)
            p2.thread_cth, p2.max_ncorrection = 85, 3
            # the count array in the REFERENCE's output order (libstdc++ unordered_map iteration order, replayed by the product's
            # host loader: dbt-hostcheck tr-order) so that .trkmc.ar compares element by element, not as multisets
            order = reference_tr_order(pref, workdir)
            ix2 = dict(ixm)
            if order is not None:
                ix2["tr_n"], ix2["tr_kmers"] = order
            ctx2 = abi.Context(p2, abi.make_index_struct(ix2))
            ctx2.run(host, offsets[: 2 * n + 1], capacity=0)
            g_tr, g_kmc, g_nmap, g_cnt = ctx2.download()
            ctx2.close()
            if thr:
                t_q, t_all, t_load, ref_counts, used = run_reference_thread(pref, host, n, workdir, nproc)
                gc = dict(zip(abi.COUNTER_NAMES, [int(x) for x in g_cnt]))
                cpu = {"value": n / t_q, "unit": "read pairs/s", "cores": used, "kind": "reference",
                       "sample": f"first {n} pairs of rank 0's batch; the reference's isThreadFeasible & helpers compiled from /root/reference "
                                 f"(oracle/_ref/ref_harness thread=85 corr={CORR} maxnc=3 aln=2, caller cpp:2072-2088 restored as glue, D1), {used} processes on "
                                 f"contiguous shards (TSV in /dev/shm); query phase {t_q:.2f}s = wall {t_all:.2f}s minus zero-read index load {t_load:.2f}s",
                       "parity_vs_gpu": {"trkmc_total_identical": int(ref_counts.sum()) == int(g_tr.sum()),
                                         "trkmc_ar_identical_ordered": bool(order is not None and np.array_equal(ref_counts, g_tr)),
                                         "trkmc_per_locus_multiset_identical": per_locus_multiset_equal(ix, ref_counts, g_tr),
                                         "gpu_sample_counters": {k: gc[k] for k in ("nThreadingReads", "nFeasibleReads", "ninc")}}}
            else:
                fa = os.path.join(workdir, "sample.fa")
                write_fasta_sample(fa, host, n)
                t_q, t_all, t_load = run_reference(pref, fa, os.path.join(workdir, "ref"), nproc)
                ref_counts = rpgg_io.read_ar(os.path.join(workdir, "ref.trkmc.ar"))
                ref_summary = open(os.path.join(workdir, "ref.tr.summary.txt")).read()
                g_summary = "".join(f"{int(x)}\t{int(y)}\n" for x, y in zip(g_nmap, g_kmc))
                cpu = {"value": n / t_q, "unit": "read pairs/s", "cores": nproc, "kind": "reference",
                       "sample": f"first {n} pairs of rank 0's batch as FASTA in /dev/shm, `danbing-tk -k {K} -kf 4 1 -cth 45 -ae -p {nproc}` (kam lines to /dev/null); "
                                 f"query phase {t_q:.2f}s = wall {t_all:.2f}s minus zero-read index load {t_load:.2f}s",
                       "parity_vs_gpu": {"tr_summary_identical": ref_summary == g_summary, "trkmc_total_identical": int(ref_counts.sum()) == int(g_tr.sum()),
                                         "trkmc_ar_identical_ordered": bool(order is not None and np.array_equal(ref_counts, g_tr)),
                                         "trkmc_per_locus_multiset_identical": per_locus_multiset_equal(ix, ref_counts, g_tr)}}
            if t_q < 0.05 * t_all:
                # the subtraction is meaningless when the index load dominates the run (20 000 loci: 44 s of load, a 0.03 s difference):
                # keep the parity check, do not quote a rate
                cpu["value"] = None
                cpu["value_note"] = "not quoted: the query phase is below 5 % of the run's wall time (index load dominates); this arm is the parity check"
        else:
            cpu = {"value": None, "unit": "read pairs/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}

    total_inc = allsum(cnt["ninc"])
    out = {"metric": METRIC, "value": value, "unit": "read pairs/s", "n_gpus": world, "steps": a.steps,
           "warmup": a.warmup, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "u64", "data": "synthetic", "config": config, "clocks": clk.summary(), "e2e": e2e, "gpu_launches": launches,
           "roofline": roofline, "cpu_baseline": cpu,
           "host_binding_rank0": dict(_numa_info, ncpus_in_affinity_mask=len(os.sched_getaffinity(0))),
           "timing": {"device_events_ms_per_step": step_ms, "host_wall_ms_per_step": wall_step_ms,
                      "note": "value uses the CUDA-event time at every N (max over ranks); at N > 1 a step ends with the host-synchronous NCCL reduce"},
           "multi_gpu_parity": mgp,
           "counters_per_step_rank0": cnt, "tr_kmer_increments_per_step_all_ranks": total_inc,
           "graph_table_bytes": int(L.dbt_graph_table_bytes(ctx.h)),
           "index_table": {"buckets": int(stats[0]), "slots_per_bucket": int(stats[1]), "keys": int(stats[2]),
                           "overflowed_buckets": int(stats[4]), "stash_keys": int(stats[5]), "bytes": int(stats[6]),
                           "overflowed_stash_buckets": int(stats[7])}}
    ctx.close()
    del host, owner, off_pin, off_owner, offsets
    return out


if __name__ == "__main__":
    main()
