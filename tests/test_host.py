"""CPU tests of the C++ host side of the drop-in CLI (danbing-tk_b200/host/*.hpp) through bin/dbt-hostcheck:
loaders + libstdc++ output-order replay, FASTA/FASTQ pairing, kam formatting, flag scanner.  No GPU needed."""
import gzip
import os
import random
import subprocess

import numpy as np
import pytest

from common import GOLDEN, golden_case, golden_outputs, golden_reads, params_for
from oracle_abi import Oracle, PKG_DIR, abi, rpgg_io

HOSTCHECK = os.path.join(PKG_DIR, "bin", "dbt-hostcheck")
pytestmark = pytest.mark.skipif(not os.path.exists(HOSTCHECK), reason="bin/dbt-hostcheck not built (run __graft_entry__.build())")


def gunzip_to(src_gz, dst):
    with gzip.open(src_gz, "rb") as f, open(dst, "wb") as o:
        o.write(f.read())
    return dst


def run(*args, **kw):
    return subprocess.run([HOSTCHECK, *map(str, args)], capture_output=True, check=False, **kw)


@pytest.mark.parametrize("case_name", ["s1", "multi"])
def test_tr_kmers_output_order_matches_reference(case_name, tmp_path):
    """read_tr_kmers_ordered reproduces the unordered_map iteration order the reference dumps in (SURVEY F4 trap 6):
    golden pan.tr.order.u64 was extracted from a reference `-on` run."""
    d = os.path.join(GOLDEN, case_name)
    tr = gunzip_to(os.path.join(d, "pan.tr.kmers.gz"), str(tmp_path / "pan.tr.kmers"))
    out = str(tmp_path / "order.u64")
    r = run("tr-order", tr, out)
    assert r.returncode == 0, r.stderr
    got = np.fromfile(out, "<u8")
    exp = np.frombuffer(gzip.open(os.path.join(d, "pan.tr.order.u64.gz")).read(), "<u8")
    assert np.array_equal(got, exp)


def test_index_loader_sizes_and_errors(tmp_path):
    d = os.path.join(GOLDEN, "multi")
    for ext in ("tr.kmers", "kmers.dbi", "fl.kdb"):
        gunzip_to(os.path.join(d, f"pan.{ext}.gz"), str(tmp_path / f"pan.{ext}"))
    r = run("index", tmp_path / "pan")
    assert r.returncode == 0, r.stderr
    ix = golden_case("multi")["ix"]
    f = dict(zip(r.stdout.split()[::2], map(int, r.stdout.split()[1::2])))
    assert f == {b"nloci": ix["nloci"], b"nk_dbi": len(ix["dbi_keys"]), b"nvv": len(ix["vv"]), b"ntr": len(ix["tr_kmers"]),
                 b"nfl": len(ix["fl_kmers"]), b"nqc": 0}
    # a missing / truncated .fl.kdb must fail loudly (the reference reads garbage, h:681)
    os.truncate(tmp_path / "pan.fl.kdb", 20)
    assert run("index", tmp_path / "pan").returncode == 1
    os.remove(tmp_path / "pan.fl.kdb")
    assert run("index", tmp_path / "pan").returncode == 1


def ref_pairing(lines, is_fastq, min_read_size, reads_per_batch, return_db=False):
    """Python restatement of the read loop of CountWords (cpp:1869-1981), incl. std::getline leaving the string
    untouched at EOF and the FASTQ branch lacking the `se and EOF` break."""
    pos = [0]
    db = {}
    out = []

    def eof():
        return pos[0] >= len(lines)

    def getline(old):
        if eof():
            return old
        pos[0] += 1
        return lines[pos[0] - 1]

    def prune(t):
        return t[:-2] if len(t) >= 2 and t[-2] == "/" and t[-1] in "12" else t
    while not eof():
        title = seq1 = seq2 = qt = qual1 = qual2 = ""
        n = 0
        while n < reads_per_batch and not eof():
            se = True
            while se:
                title = getline(title)
                seq1 = getline(seq1)
                if is_fastq:
                    qt = getline(qt)
                    qual1 = getline(qual1)
                title = prune(title)
                if title in db:
                    s2, q2 = db.pop(title)
                    if len(seq1) < min_read_size or len(s2) < min_read_size:
                        continue
                    seq2, qual2 = s2, q2
                    se = False
                    break
                db[title] = (seq1, qual1)
                if eof():
                    break
            if not is_fastq and se and eof():
                break
            out.append((title, seq1, seq2, qual1, qual2))
            n += 2
    return (out, db) if return_db else (out, len(db))


def make_fastx(rng, npairs, is_fastq, shuffle, singletons):
    recs = []
    for i in range(npairs):
        la, lb = rng.choice([150, 150, 150, 90, 40, 20, 151]), rng.choice([150, 150, 120, 30])
        a = "".join(rng.choice("ACGTN") for _ in range(la))
        b = "".join(rng.choice("ACGT") for _ in range(lb))
        recs.append((f"r{i}/1", a))
        recs.append((f"r{i}/2", b))
    for i in range(singletons):
        recs.append((f"single{i}", "ACGT" * 30))
    if shuffle:
        # move some mates far away from their partner
        for _ in range(npairs // 3):
            i, j = rng.randrange(len(recs)), rng.randrange(len(recs))
            recs[i], recs[j] = recs[j], recs[i]
    lines = []
    for t, s in recs:
        if is_fastq:
            lines += ["@" + t, s, "+", "".join(chr(33 + rng.randrange(40)) for _ in s)]
        else:
            lines += [">" + t, s]
    return lines


@pytest.mark.parametrize("is_fastq", [0, 1])
@pytest.mark.parametrize("shuffle,singletons,trailing_newline", [(False, 0, True), (True, 3, True), (True, 1, False), (False, 1, True)])
def test_reader_pairs_like_the_reference(tmp_path, is_fastq, shuffle, singletons, trailing_newline):
    rng = random.Random(7 * is_fastq + 3 * shuffle + singletons)
    lines = make_fastx(rng, 400, is_fastq, shuffle, singletons)
    if singletons and not shuffle:
        lines = lines + lines[:0]        # singleton is the LAST record: exercises the EOF handling of both branches
    fn = str(tmp_path / "reads.fx")
    with open(fn, "w") as f:
        f.write("\n".join(lines) + ("\n" if trailing_newline else ""))
    for ref_batch, cap in ((300000, 1 << 12), (64, 1 << 12), (64, 50)):
        exp, ndb = ref_pairing(lines, bool(is_fastq), 55, ref_batch)
        r = run("pair", fn, is_fastq, 55, ref_batch, cap)
        assert r.returncode == 0, r.stderr
        got = [tuple(l.split("\t")) for l in r.stdout.decode().split("\n")[:-1]]
        want = [(t, s1, s2) + ((q1, q2) if is_fastq else ()) for t, s1, s2, q1, q2 in exp]
        assert got == want, (ref_batch, cap)
        assert r.stderr.decode().split() == ["reads", str(2 * len(exp)), "unpaired", str(ndb)]


@pytest.mark.parametrize("case_name,run_name", [("s1", "cth45"), ("multi", "cth10"), ("multi", "kf0")])
def test_kam_writer_matches_reference_stdout(case_name, run_name, tmp_path):
    """records of the CPU oracle -> C++ kam formatter == the reference binary's stdout (golden out.kam)"""
    case = golden_case(case_name)
    runp = case["meta"]["runs"][run_name]
    titles, seqs = golden_reads(case, run_name)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    p = params_for(runp, flags=abi.F_KAM | abi.F_TRACE)
    recs, cnt, tr, kmc, nmap = Oracle(case["ix"]).run(p, bases, off)
    keep = recs[recs["h"]["status"] == 6]
    recfn = str(tmp_path / "recs.bin")
    keep.tofile(recfn)
    fa = gunzip_to(case["fasta"] + ".gz", str(tmp_path / "reads.fa")) if not os.path.exists(case["fasta"]) else case["fasta"]
    r = run("kam", recfn, keep.dtype.itemsize, keep.dtype["as"].shape[1], fa, 0, runp["cth"] + runp["k"] - 1)
    assert r.returncode == 0, r.stderr
    assert r.stdout.decode().splitlines() == golden_outputs(case, run_name)["kam"]


def test_flag_scanner(tmp_path):
    for n in ("p.tr.kmers", "r.fa", "q.txt"):
        open(tmp_path / n, "w").close()
    pref, fa, out = str(tmp_path / "p"), str(tmp_path / "r.fa"), str(tmp_path / "o")
    r = run("args", "-k", 25, "-gc", 85, 3, "-ae", "-kf", 4, 1, "-cth", 45, "-qc", tmp_path / "q.txt", "-qs", pref, "-fa", fa, "-o", out, "-p", 4)
    assert r.returncode == 0, r.stdout
    s = r.stdout.decode()
    assert "k=25 kf=4,1 cth=45 c=40 p=4 r=1 threading=1 thread_cth=85 maxnc=3 tc=0 aln=1 alnmin=1 okam=1 qc=1 fq=0 named=0" in s
    assert os.path.exists(out + ".trkmc.ar")                    # created while parsing, like cpp:2417-2419
    assert b"graph threading mode: 1" in r.stderr and b"Cthreshold: 45" in r.stderr
    # -gc INT1 without INT2 keeps the default edit budget; -on; -ka; -fq
    r = run("args", "-gc", 100, "-ka", "-qs", pref, "-fq", fa, "-on", out)
    assert "thread_cth=100 maxnc=4" in r.stdout.decode() and "okam=0" in r.stdout.decode() and "fq=1 named=1" in r.stdout.decode()
    assert run("args", "-zz").stdout.startswith(b"error: invalid option: -zz")
    assert run("args", "-qs", str(tmp_path / "missing")).returncode == 1


# ---- raw ingestion: RawReader + StreamPairer around the GPU's neighbour pairing ----
def run_pair_raw(tmp_path, lines, is_fastq, ref_batch, target, trailing_newline=True, minlen=55):
    fn = str(tmp_path / "reads.fx")
    with open(fn, "w") as f:
        f.write("\n".join(lines) + ("\n" if trailing_newline and lines else ""))
    exp, ndb = ref_pairing(lines, bool(is_fastq), minlen, ref_batch)
    r = run("pair-raw", fn, int(is_fastq), minlen, ref_batch, target)
    assert r.returncode == 0, r.stderr
    got = [tuple(l.split("\t")) for l in r.stdout.decode().split("\n")[:-1]]
    want = [(t, s1, s2) + ((q1, q2) if is_fastq else ()) for t, s1, s2, q1, q2 in exp]
    assert got == want, (ref_batch, target)
    st = r.stderr.decode().split()
    assert st[:4] == ["reads", str(2 * len(exp)), "unpaired", str(ndb)]
    return dict(fast=int(st[5]), slow=int(st[7]))


@pytest.mark.parametrize("is_fastq", [0, 1])
@pytest.mark.parametrize("shuffle,singletons,trailing_newline", [(False, 0, True), (True, 3, True), (True, 1, False), (False, 1, True)])
def test_raw_reader_and_pairer_like_the_reference(tmp_path, is_fastq, shuffle, singletons, trailing_newline):
    """same inputs as the line reader's test, cut into chunks of 64 KB / 3 KB / 700 B so that mates, stragglers and the
    end-of-file rules meet chunk borders"""
    rng = random.Random(7 * is_fastq + 3 * shuffle + singletons)
    lines = make_fastx(rng, 400, is_fastq, shuffle, singletons)
    for ref_batch, target in ((300000, 1 << 16), (64, 3000), (64, 700), (2, 1200)):
        st = run_pair_raw(tmp_path, lines, is_fastq, ref_batch, target, trailing_newline)
        if not shuffle and not singletons:
            assert st["slow"] == 0          # name-collated input never leaves the GPU path
        # no mate is "short" at this threshold: stragglers are then paired by the host map beside the GPU's neighbour pairs
        st = run_pair_raw(tmp_path, lines, is_fastq, ref_batch, target, trailing_newline, minlen=10)
        assert st["slow"] == 0 and st["fast"] >= 1


@pytest.mark.parametrize("is_fastq", [0, 1])
def test_raw_pairer_runs_of_equal_titles_and_third_mates(tmp_path, is_fastq):
    """three and four records with one title, a neighbour pair whose title a straggler is waiting for, duplicated pairs:
    the order of arrival decides in the reference's map -- such chunks must come back from the host pairer identical"""
    rng = random.Random(99 + is_fastq)
    lines = make_fastx(rng, 120, is_fastq, False, 0)
    L = 4 if is_fastq else 2
    recs = [lines[i:i + L] for i in range(0, len(lines), L)]

    def retitle(r, t):
        return [r[0][0] + t] + r[1:]
    out = []
    for i in range(0, len(recs), 2):
        a, b = recs[i], recs[i + 1]
        if i % 40 == 6:
            out += [a, b, retitle(a, b[0][1:-2] + "/1")]                     # three in a row
        elif i % 40 == 14:
            out += [a, b, retitle(a, b[0][1:-2] + "/1"), retitle(b, b[0][1:-2] + "/2")]   # four in a row
        elif i % 40 == 22:
            out += [retitle(a, "late%d/1" % i)] + recs[i + 2:i + 4] + [retitle(a, "late%d/2" % i), retitle(b, "late%d/1" % i)]   # straggler, then a neighbour pair of its title
        else:
            out += [a, b]
    flat = [l for r in out for l in r]
    for ref_batch, target in ((300000, 1 << 16), (10, 2000), (4, 900)):
        st = run_pair_raw(tmp_path, flat, is_fastq, ref_batch, target)
        assert st["slow"] >= 1


@pytest.mark.parametrize("is_fastq", [0, 1])
@pytest.mark.parametrize("drop", [1, 2, 3])
def test_raw_reader_truncated_last_record_and_empty_input(tmp_path, is_fastq, drop):
    """a file that ends in the middle of a record: std::getline leaves the reference's strings alone at end of file"""
    rng = random.Random(5 + drop)
    lines = make_fastx(rng, 30, is_fastq, False, 0)
    L = 4 if is_fastq else 2
    if drop >= L:
        pytest.skip("a FASTA record has two lines")
    for ref_batch, target in ((300000, 1 << 16), (6, 800)):
        run_pair_raw(tmp_path, lines[:-drop], is_fastq, ref_batch, target)
        run_pair_raw(tmp_path, lines[:L * 7 - drop], is_fastq, ref_batch, target)     # the truncated record's mate is missing
    run_pair_raw(tmp_path, [], is_fastq, 300000, 1 << 16)
    run_pair_raw(tmp_path, lines[:1], is_fastq, 300000, 1 << 16, trailing_newline=False)


def test_raw_reader_parallel_pread(tmp_path):
    """a regular file is read by several threads at once (pread): same chunks, same pairs"""
    rng = random.Random(31)
    lines = make_fastx(rng, 30000, 0, False, 0)
    run_pair_raw(tmp_path, lines, 0, 300000, 5 << 20)


def test_raw_reader_inflates_gzip(tmp_path):
    """a .gz input (magic sniffed, or named *.gz) gives the same pairs as the plain text"""
    rng = random.Random(77)
    lines = make_fastx(rng, 3000, 1, True, 2)
    exp, ndb = ref_pairing(lines, True, 55, 300000)
    for name in ("reads.fq.gz", "reads.bin"):
        fn = str(tmp_path / name)
        with gzip.open(fn, "wt") as f:
            f.write("\n".join(lines) + "\n")
        r = run("pair-raw", fn, 1, 55, 300000, 40000)
        assert r.returncode == 0, r.stderr
        got = [tuple(l.split("\t")) for l in r.stdout.decode().split("\n")[:-1]]
        assert got == [(t, s1, s2, q1, q2) for t, s1, s2, q1, q2 in exp]
        assert r.stderr.decode().split()[:4] == ["reads", str(2 * len(exp)), "unpaired", str(ndb)]


def ref_qmask(qual, qth, k):
    """qString2qMask restated (src/aQueryFasta_thread.h:1038-1071): the scan's base index is bounded by the number of k-mers"""
    nq = len(qual)
    nk = nq - k + 1
    qs = [ord(c) - 33 for c in qual]
    m = [False] * max(nk, 0)
    if nk < 1:
        return m
    qi = ki = 0
    while qs[qi] < qth:
        qi += 1
        ki += 1
        if qi >= nk:
            return m
    while qi < nk:
        ok = True
        qj = qi
        while qi < qj + k:
            if qs[qi] < qth:
                ok = False
                ki = qi
                while qs[qi] < qth:
                    qi += 1
                    ki += 1
                    if qi >= nk:
                        return m
                break
            qi += 1
        if ok:
            m[ki] = True
            ki += 1
            if qi >= nk:
                return m
            while qs[qi] >= qth:
                m[ki] = True
                qi += 1
                ki += 1
                if qi >= nk:
                    return m
            ki = qi
            while qs[qi] < qth:
                qi += 1
                ki += 1
                if qi >= nk:
                    return m
    return m


def test_quality_mask_matches_the_reference_scan():
    """`-b` with FASTQ: host/bait.hpp quality_mask (the device's quality_mask in csrc/bait.cuh is the same state machine, checked
    against the live reference in tests/test_cli_gpu.py) against the restated qString2qMask, incl. its early stop k-1 bases
    before the end of the read"""
    rng = random.Random(5)
    quals = []
    for n in [21, 22, 40, 41, 42, 63, 100, 150, 151, 250]:
        for _ in range(40):
            q = [rng.choice("FGHI5") for _ in range(n)]
            for _ in range(rng.randrange(5)):
                o, w = rng.randrange(n), rng.randrange(1, 25)
                q[o:o + w] = [rng.choice("#,+4") for _ in q[o:o + w]]
            quals.append("".join(q))
        quals += ["#" * n, "I" * n, "#" + "I" * (n - 1), "I" * (n - 1) + "#"]
    r = subprocess.run([HOSTCHECK, "qmask", "21", "20"], input="\n".join(quals) + "\n", capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = r.stdout.split("\n")[:-1]
    assert len(got) == len(quals)
    nset = 0
    for q, g in zip(quals, got):
        e = "".join("1" if x else "0" for x in ref_qmask(q, 20, 21))
        assert g == e, q
        nset += e.count("1")
    assert nset > 1000
    # the all-high read of 150 bases (130 k-mers): the scan stops k-1 bases early, so the last k-1 k-mers stay out
    assert got[quals.index("I" * 150)].count("1") == 110
