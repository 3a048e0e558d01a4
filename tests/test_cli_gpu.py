"""GPU tests of the drop-in CLI (danbing-tk_b200/bin/danbing-tk): byte-for-byte against the golden outputs of the
unmodified reference binary, and -- where oracle/_ref travelled to the box -- against the reference run live."""
import gzip
import os
import random
import re
import subprocess

import numpy as np
import pytest

from common import GOLDEN, golden_case, golden_outputs
from oracle_abi import PKG_DIR, abi, ref_bin

pytestmark = pytest.mark.gpu
CLI = os.path.join(PKG_DIR, "bin", "danbing-tk")
COUNTER_RE = re.compile(r"^(\d+) reads (.*)\.$", re.M)


def stage_case(case_name, tmp_path):
    d = os.path.join(GOLDEN, case_name)
    for ext in ("tr.kmers", "kmers.dbi", "fl.kdb", "tre.kdb"):
        with gzip.open(os.path.join(d, f"pan.{ext}.gz"), "rb") as f, open(tmp_path / f"pan.{ext}", "wb") as o:
            o.write(f.read())
    with gzip.open(os.path.join(d, "reads.fa.gz"), "rb") as f, open(tmp_path / "reads.fa", "wb") as o:
        o.write(f.read())
    return str(tmp_path / "pan"), str(tmp_path / "reads.fa")


def run_cli(exe, args, cwd):
    r = subprocess.run([exe, *map(str, args)], capture_output=True, cwd=cwd)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    return r.stdout, r.stderr.decode()


def counters(stderr):
    return {name: int(v) for v, name in COUNTER_RE.findall(stderr)}


def flags_of(run):
    return ["-k", run["k"], "-kf", run["kf"][0], run["kf"][1], "-cth", run["cth"], "-c", run["c"]]


READERS = [pytest.param([], id="gpu-reader"), pytest.param(["--host-reader"], id="host-reader")]   # raw text split on the GPU (default) / line reader on the host


@pytest.mark.parametrize("reader", READERS)
@pytest.mark.parametrize("case_name,run_name", [("s1", "cth45"), ("s1", "cth10"), ("multi", "cth45"), ("multi", "kf0")])
def test_cli_matches_golden(case_name, run_name, reader, tmp_path):
    case = golden_case(case_name)
    run = case["meta"]["runs"][run_name]
    pref, fa = stage_case(case_name, tmp_path)
    out, err = run_cli(CLI, flags_of(run) + reader + ["-qs", pref, "-fa", fa, "-o", tmp_path / "out"], tmp_path)
    exp = golden_outputs(case, run_name)
    assert out.decode().splitlines() == exp["kam"]
    assert np.array_equal(np.fromfile(tmp_path / "out.trkmc.ar", "<u8")[1:], exp["trkmc"])
    assert open(tmp_path / "out.tr.summary.txt").read() == exp["summary"]
    c = counters(err)
    for name, v in run["counters"].items():
        assert c[name[6:] if name.startswith("reads ") else name] == v, name


@pytest.mark.parametrize("reader", READERS)
def test_cli_small_batches_two_slots(reader, tmp_path):
    """many GPU submissions (reference batches of 600 reads, 512 pairs per submission): same bytes"""
    case = golden_case("multi")
    run = case["meta"]["runs"]["cth45"]
    pref, fa = stage_case("multi", tmp_path)
    out, err = run_cli(CLI, flags_of(run) + reader + ["-r", 0.002, "--batch-pairs", 512, "-qs", pref, "-fa", fa, "-o", tmp_path / "out"], tmp_path)
    exp = golden_outputs(case, "cth45")
    assert out.decode().splitlines() == exp["kam"]
    assert np.array_equal(np.fromfile(tmp_path / "out.trkmc.ar", "<u8")[1:], exp["trkmc"])
    assert err.count("Batch query in") > 4


needs_ref = pytest.mark.skipif(not ref_bin("danbing-tk"), reason="oracle/_ref/danbing-tk not present")


def both(args, tmp_path, extra=()):
    a = run_cli(CLI, args + list(extra) + ["-o", tmp_path / "gpu"], tmp_path)
    b = run_cli(ref_bin("danbing-tk"), args + ["-o", tmp_path / "ref"], tmp_path)
    return a, b


@needs_ref
@pytest.mark.parametrize("extra", [pytest.param([], id="gpu-reader"), pytest.param(["--batch-pairs", 300], id="gpu-reader-small-chunks"),
                                   pytest.param(["--host-reader"], id="host-reader")])
def test_cli_vs_reference_live_fastq_qc_and_stragglers(extra, tmp_path):
    """FASTQ input with shuffled mates, short mates, a trailing singleton (the reference pairs it with a stale mate,
    cpp:1919-1948), -qc mask: stdout, .trkmc.ar, .tr.summary.txt and the final counters of both binaries"""
    pref, fa = stage_case("multi", tmp_path)
    rng = random.Random(5)
    lines = open(fa).read().split("\n")[:-1]
    recs = [(lines[i], lines[i + 1]) for i in range(0, len(lines), 2)]
    for _ in range(len(recs) // 5):
        i, j = rng.randrange(len(recs)), rng.randrange(len(recs))
        recs[i], recs[j] = recs[j], recs[i]
    recs.append((">lonely/1", recs[0][1]))
    fq = tmp_path / "reads.fq"
    with open(fq, "w") as f:
        for t, s in recs:
            f.write("@" + t[1:] + "\n" + s + "\n+\n" + "".join(chr(33 + rng.randrange(41)) for _ in s) + "\n")
    nloci = golden_case("multi")["ix"]["nloci"]
    with open(tmp_path / "qc.txt", "w") as f:
        f.write("".join("0" if l % 4 == 1 else "1" for l in range(nloci)) + "\n")
    args = ["-k", 21, "-kf", 4, 1, "-cth", 30, "-qc", tmp_path / "qc.txt", "-qs", pref, "-fq", fq]
    (go, ge), (ro, re_) = both(args, tmp_path, extra)
    assert go == ro
    assert open(tmp_path / "gpu.trkmc.ar", "rb").read() == open(tmp_path / "ref.trkmc.ar", "rb").read()
    assert open(tmp_path / "gpu.tr.summary.txt").read() == open(tmp_path / "ref.tr.summary.txt").read()
    assert counters(ge) == counters(re_) and counters(ge)["removed by QC filter"] > 0


@needs_ref
@pytest.mark.parametrize("is_fastq", [False, True])
def test_cli_vs_reference_live_collated_reads_on_the_gpu_reader(is_fastq, tmp_path):
    """name-collated input (mates are neighbours) in many small chunks: every pair is formed on the GPU; a run of three
    equal titles, a short last pair and a chunk border inside a pair must not change a byte"""
    pref, fa = stage_case("multi", tmp_path)
    lines = open(fa).read().split("\n")[:-1]
    recs = [(lines[i], lines[i + 1]) for i in range(0, len(lines), 2)]
    recs.insert(40, (recs[38][0][:-2] + "/1", recs[38][1]))                   # a third record of one title
    recs += [(">tail/1", recs[0][1]), (">tail/2", recs[1][1][:20])]          # the stream ends on a pair with a short mate
    rng = random.Random(3)
    fx = tmp_path / ("reads.fq" if is_fastq else "reads2.fa")
    with open(fx, "w") as f:
        for t, s in recs:
            if is_fastq:
                f.write("@" + t[1:] + "\n" + s + "\n+\n" + "".join(chr(33 + rng.randrange(41)) for _ in s) + "\n")
            else:
                f.write(t + "\n" + s + "\n")
    args = ["-k", 21, "-kf", 4, 1, "-cth", 30, "-qs", pref, "-fq" if is_fastq else "-fa", fx]
    for extra in (["--batch-pairs", 256, "--timing"], []):
        (go, ge), (ro, re_) = both(args, tmp_path, extra)
        assert go == ro
        assert open(tmp_path / "gpu.trkmc.ar", "rb").read() == open(tmp_path / "ref.trkmc.ar", "rb").read()
        assert open(tmp_path / "gpu.tr.summary.txt").read() == open(tmp_path / "ref.tr.summary.txt").read()
        assert counters(ge) == counters(re_)
        if "--timing" in extra:
            m = re.search(r"chunks (\d+) \((\d+) paired on the host\)", ge)
            assert m and int(m.group(1)) > 4 and 1 <= int(m.group(2)) < int(m.group(1))   # only the chunk with the run of three left the GPU path


@needs_ref
def test_cli_vs_reference_live_named_output_and_threading_flag(tmp_path):
    """-on writes OUT.tr.kmers (k-mer names in the reference's hashtable order); -gc at HEAD counts nothing"""
    pref, fa = stage_case("s1", tmp_path)
    for exe, tag in ((CLI, "gpu"), (ref_bin("danbing-tk"), "ref")):
        run_cli(exe, ["-k", 21, "-cth", 45, "-ka", "-qs", pref, "-fa", fa, "-on", tmp_path / tag], tmp_path)
    assert open(tmp_path / "gpu.tr.kmers").read() == open(tmp_path / "ref.tr.kmers").read()
    (go, ge), (ro, re_) = both(["-k", 21, "-gc", 85, 3, "-ae", "-kf", 4, 1, "-cth", 45, "-qs", pref, "-fa", fa], tmp_path)
    assert go == ro == b""
    assert open(tmp_path / "gpu.trkmc.ar", "rb").read() == open(tmp_path / "ref.trkmc.ar", "rb").read()
    assert counters(ge) == counters(re_) and counters(ge)["passsed threading"] == 0 and counters(ge)["entered threading step"] > 0


def test_cli_two_gpus_same_bytes(tmp_path, lib):
    if lib.dbt_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    case = golden_case("multi")
    run = case["meta"]["runs"]["cth45"]
    pref, fa = stage_case("multi", tmp_path)
    out, err = run_cli(CLI, flags_of(run) + ["-r", 0.002, "--batch-pairs", 512, "--gpus", 2, "-qs", pref, "-fa", fa, "-o", tmp_path / "out"], tmp_path)
    exp = golden_outputs(case, "cth45")
    assert out.decode().splitlines() == exp["kam"]
    assert np.array_equal(np.fromfile(tmp_path / "out.trkmc.ar", "<u8")[1:], exp["trkmc"])
    assert open(tmp_path / "out.tr.summary.txt").read() == exp["summary"]


# ---- restored graph threading (-g / -gc with PREF.graph.kmers): .aln lines and counts vs the reference harness golden ----
def stage_thread_case(tmp_path):
    import json
    import thread_common as tc
    from test_oracle_thread import GOLD, split_case
    meta = json.load(open(os.path.join(GOLD, "meta.json")))
    gk, pk = split_case(meta["case"])
    pref, ix, seqs, _ = tc.thread_case(outdir=str(tmp_path), **gk)
    fa = str(tmp_path / "reads.fa")
    with open(fa, "w") as f:
        for i in range(len(seqs) // 2):
            f.write(f">p{i}/1\n{seqs[2 * i + 1]}\n>p{i}/2\n{seqs[2 * i]}\n")
    flags = ["-k", 21, "-kf", 4, 1, "-cth", pk["cth"], "-gc", pk["thread"], pk["maxnc"]]
    return pref, fa, flags, meta, GOLD


@pytest.mark.parametrize("mode,gold", [("-a", "out.aln.gz"), ("-ae", "out.aln_min.gz")])
def test_cli_threading_aln_lines_and_counts_match_reference_harness(mode, gold, tmp_path):
    pref, fa, flags, meta, GOLD = stage_thread_case(tmp_path)
    out, err = run_cli(CLI, flags + [mode, "-qs", pref, "-fa", fa, "-o", tmp_path / "out"], tmp_path)
    exp = gzip.open(os.path.join(GOLD, gold), "rt").read()
    assert out.decode() == exp
    gold_tr = np.frombuffer(gzip.open(os.path.join(GOLD, "out.trkmc.ar.gz")).read(), "<u8")
    assert np.array_equal(np.fromfile(tmp_path / "out.trkmc.ar", "<u8"), gold_tr)
    c = counters(err)
    assert c["entered threading step"] == meta["counters"]["nThreadingReads"]
    assert c["passsed threading"] == meta["counters"]["nFeasibleReads"]
    assert "graph threading restored" in err


def test_cli_threading_without_alignment_output_and_legacy_graph_umap(tmp_path):
    """no -a/-ae: empty stdout, same counts; the graph may also come as a v1.3 PREF.graph.umap"""
    pref, fa, flags, meta, GOLD = stage_thread_case(tmp_path)
    from oracle_abi import rpgg_io
    n, k, v = rpgg_io.read_kmers_text(pref + ".graph.kmers")
    with open(pref + ".graph.umap", "wb") as f:
        f.write(np.uint64(len(n)).tobytes())
        o = 0
        for c in n:
            c = int(c)
            f.write(np.uint64(c).tobytes())
            rec = np.zeros(c, dtype=[("k", "<u8"), ("m", "u1")])
            rec["k"], rec["m"] = k[o:o + c], v[o:o + c].astype("u1")
            f.write(rec.tobytes())
            o += c
    os.remove(pref + ".graph.kmers")
    out, err = run_cli(CLI, flags + ["-qs", pref, "-fa", fa, "-o", tmp_path / "out"], tmp_path)
    assert out == b""
    gold_tr = np.frombuffer(gzip.open(os.path.join(GOLD, "out.trkmc.ar.gz")).read(), "<u8")
    assert np.array_equal(np.fromfile(tmp_path / "out.trkmc.ar", "<u8"), gold_tr)


# ---- the remaining align flags, byte for byte against the live reference ----
def retitled(fa, tmp_path, title_of, is_fastq=False, name="reads_t"):
    lines = open(fa).read().split("\n")[:-1]
    rng = random.Random(17)
    fx = tmp_path / (name + (".fq" if is_fastq else ".fa"))
    with open(fx, "w") as f:
        for i in range(0, len(lines), 4):                          # two records = one pair
            t = title_of(i // 4)
            for m, (s, tag) in enumerate(((lines[i + 1], "/1"), (lines[i + 3], "/2"))):
                if is_fastq:
                    f.write("@" + t + tag + "\n" + s + "\n+\n" + "".join(chr(33 + rng.randrange(41)) for _ in s) + "\n")
                else:
                    f.write(">" + t + tag + "\n" + s + "\n")
    return fx


@needs_ref
@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("is_fastq", [False, True])
def test_cli_vs_reference_live_extract_reads(mode, is_fastq, tmp_path):
    """`-e 1|2` (cpp:2094-2099, 1618-1644): pairs that pass the locus vote are written out as FASTA/FASTQ, nothing is counted"""
    pref, fa = stage_case("multi", tmp_path)
    fx = retitled(fa, tmp_path, lambda i: f"frag{i}", is_fastq)
    args = ["-k", 21, "-kf", 4, 1, "-cth", 30, "-e", mode, "-qs", pref, "-fq" if is_fastq else "-fa", fx]
    for extra in ([], ["--batch-pairs", 300], ["--host-reader"]):
        (go, ge), (ro, re_) = both(args, tmp_path, extra)
        assert go == ro and len(go) > 10000
        assert counters(ge) == counters(re_) and counters(ge)["passsed threading"] > 0
        assert not os.path.exists(tmp_path / "gpu.tr.summary.txt")


@needs_ref
@pytest.mark.parametrize("simmode", [1, 2])
def test_cli_vs_reference_live_simulation_truth_in_kam(simmode, tmp_path):
    """`-s 1|2` (cpp:476-524, 2169): the source locus read off the read name fills the first kam column, and pairs that
    assignTRkmc removed are printed too when their source locus is known"""
    pref, fa = stage_case("multi", tmp_path)
    nloci = golden_case("multi")["ix"]["nloci"]
    if simmode == 1:
        title_of = lambda i: f"{(i // 7) % nloci}.frag{i}"
    else:
        title_of = lambda i: f"chr1:{i}-{i + 150}:" + ("." if (i // 5) % 4 == 3 else str((i // 5) % nloci))
    fx = retitled(fa, tmp_path, title_of)
    args = ["-k", 21, "-kf", 4, 1, "-cth", 30, "-c", 1000, "-s", simmode, "-qs", pref, "-fa", fx]   # -c 1000: spanning reads are removed by assignTRkmc
    for extra in ([], ["--batch-pairs", 300], ["--host-reader"]):
        (go, ge), (ro, re_) = both(args, tmp_path, extra)
        assert go == ro and len(go) > 10000
        assert open(tmp_path / "gpu.trkmc.ar", "rb").read() == open(tmp_path / "ref.trkmc.ar", "rb").read()
        assert counters(ge) == counters(re_)
    removed = [l for l in go.decode().splitlines() if l.split("\t")[1] == str(nloci)]
    assert removed, "no removed pair with a known source locus in this input: the test does not exercise cpp:2169"


@needs_ref
@pytest.mark.parametrize("extra", [pytest.param([], id="gpu-reader"), pytest.param(["--batch-pairs", 300], id="gpu-reader-small-chunks"),
                                   pytest.param(["--host-reader", "-r", 0.004], id="host-reader-small-ref-batches")])
def test_cli_vs_reference_live_bubbles(extra, tmp_path):
    """`-bu` (cpp:1559-1606): reads that carry a recurring variant absent from the graph -> OUT.bub.kmdb (edges seen >= 5 times, in the
    reference's hashtable order) byte for byte; the reference runs with -p 1 (its batches merge in completion order otherwise)"""
    pref, fa = stage_case("multi", tmp_path)
    lines = open(fa).read().split("\n")[:-1]
    seqs = lines[1::2]
    rng = random.Random(23)
    comp = {"A": "C", "C": "G", "G": "T", "T": "A"}
    for _ in range(150):                                   # a variant = the middle base of a 31-mer of some read, changed in every read that has it
        s = seqs[rng.randrange(len(seqs))]
        if len(s) < 80:
            continue
        o = rng.randrange(20, len(s) - 51)
        seed = s[o:o + 31]
        var = seed[:15] + comp.get(seed[15], "A") + seed[16:]
        seqs = [x.replace(seed, var) for x in seqs]
    fx = tmp_path / "reads_var.fa"
    with open(fx, "w") as f:
        for t, s in zip(lines[0::2], seqs):
            f.write(t + "\n" + s + "\n")
    rflag = [x for x in extra if x in ("-r", 0.004)]
    args = ["-k", 21, "-kf", 4, 1, "-cth", 30, "-bu", "-p", 1] + rflag + ["-qs", pref, "-fa", fx]
    a = run_cli(CLI, args + [x for x in extra if x not in rflag] + ["-o", tmp_path / "gpu"], tmp_path)
    b = run_cli(ref_bin("danbing-tk"), args + ["-o", tmp_path / "ref"], tmp_path)
    assert a[0] == b[0]
    assert open(tmp_path / "gpu.trkmc.ar", "rb").read() == open(tmp_path / "ref.trkmc.ar", "rb").read()
    bub = open(tmp_path / "ref.bub.kmdb", "rb").read()
    assert open(tmp_path / "gpu.bub.kmdb", "rb").read() == bub
    nloci = golden_case("multi")["ix"]["nloci"]
    nk = int(np.frombuffer(bub, "<u8", 1, 8 + 8 * nloci)[0])
    assert nk >= 10, "the input produced too few recurring novel edges to test anything"
    assert counters(a[1]) == counters(b[1])


@needs_ref
@pytest.mark.parametrize("is_fastq", [False, True], ids=["fa", "fq"])
@pytest.mark.parametrize("track", [False, True], ids=["b", "b-tb"])
def test_cli_vs_reference_live_bait_filter(track, is_fastq, tmp_path):
    """`-b` (cpp:1377-1419, 2110-2125) and `-tb`: a bait file built by the reference's own `ktools serialize-bt`; pairs whose
    mates hold a bait k-mer of the voted locus outside its [min, max] occurrence bounds are removed.  With FASTQ only k-mers whose
    bases all reach `-qth` count (qString2qMask, h:1038-1071; qualities here have low runs at the ends and inside the reads).  stdout,
    count files, counters and (with -tb) OUT.btk.kmdb must equal the reference's (-p 1: its tracker merges batches in completion
    order otherwise)."""
    pref, fa = stage_case("multi", tmp_path)
    if is_fastq:
        qrng = random.Random(77)

        def quality(n):
            q = [qrng.choice("FGHI:5") for _ in range(n)]
            for _ in range(qrng.randrange(4)):                 # a few low-quality stretches, anywhere incl. both ends
                o, w = qrng.randrange(n), qrng.randrange(1, 12)
                q[o:o + w] = [qrng.choice("#,+4") for _ in q[o:o + w]]
            if qrng.random() < 0.02:
                q = ["#"] * n
            return "".join(q)
        fq = tmp_path / "reads_b.fq"
        lines = open(fa).read().split("\n")[:-1]
        with open(fq, "w") as f:
            for t, sq in zip(lines[0::2], lines[1::2]):
                f.write("@" + t[1:] + "\n" + sq + "\n+\n" + quality(len(sq)) + "\n")
        fa = fq
    nloci = golden_case("multi")["ix"]["nloci"]
    rng = random.Random(31)
    per, l = [[] for _ in range(nloci)], -1
    for line in open(pref + ".tr.kmers"):
        if line[0] == ">":
            l += 1
        else:
            per[l].append(int(line.split()[0]))
    with open(tmp_path / "bait.txt", "w") as f:
        for l in range(nloci):
            f.write(f">{l}\n")
            for km in rng.sample(per[l], min(len(per[l]), 40)):
                mi, ma = rng.choice([(0, 0), (2, 255), (1, 255), (255, 0), (1, 1), (0, 3)])
                f.write(f"{km}\t{mi}\t{ma}\n")
    r = subprocess.run([ref_bin("ktools"), "serialize-bt", str(tmp_path / "bait.txt"), str(nloci), pref], capture_output=True)
    assert r.returncode == 0, r.stderr
    args = ["-k", 21, "-kf", 4, 1, "-cth", 30, "-b"] + (["-tb"] if track else []) + (["-qth", 21] if is_fastq else []) + ["-p", 1, "-qs", pref, "-fq" if is_fastq else "-fa", fa]
    for extra in ([], ["--batch-pairs", 300], ["--host-reader", "-r", 0.004]):
        rflag = [x for x in extra if x in ("-r", 0.004)]
        a = run_cli(CLI, args[:-4] + rflag + args[-4:] + [x for x in extra if x not in rflag] + ["-o", tmp_path / "gpu"], tmp_path)
        b = run_cli(ref_bin("danbing-tk"), args[:-4] + rflag + args[-4:] + ["-o", tmp_path / "ref"], tmp_path)
        assert a[0] == b[0]
        assert open(tmp_path / "gpu.trkmc.ar", "rb").read() == open(tmp_path / "ref.trkmc.ar", "rb").read()
        assert open(tmp_path / "gpu.tr.summary.txt").read() == open(tmp_path / "ref.tr.summary.txt").read()
        ca, cb = counters(a[1]), counters(b[1])
        assert ca == cb and ca["removed by bait locus"] > 20
        if track:
            btk = open(tmp_path / "ref.btk.kmdb", "rb").read()
            assert open(tmp_path / "gpu.btk.kmdb", "rb").read() == btk
            assert int(np.frombuffer(btk, "<u8", 1, 8 + 8 * nloci)[0]) >= 5
    if is_fastq and not track:
        # one phred score per base: a record whose quality line is shorter than its sequence is refused by name, with either reader
        lines = open(fa).read().split("\n")
        lines[4 * 777 + 3] = lines[4 * 777 + 3][:-7]
        bad = tmp_path / "reads_badq.fq"
        open(bad, "w").write("\n".join(lines))
        for extra in ([], ["--host-reader"]):
            r = subprocess.run([CLI, *map(str, args[:-1] + [bad] + extra + ["-o", tmp_path / "bad"])], capture_output=True, cwd=tmp_path)
            assert r.returncode != 0 and b"quality string is not as long as its sequence" in r.stderr, r.stderr[-500:]


def test_cli_gzip_input_and_stdin(tmp_path):
    """the GPU reader takes a gzip file (inflated on the fly) and a pipe on /dev/stdin like a plain file: same bytes"""
    case = golden_case("multi")
    run = case["meta"]["runs"]["cth45"]
    pref, fa = stage_case("multi", tmp_path)
    exp = golden_outputs(case, "cth45")
    with open(fa, "rb") as f, gzip.open(str(tmp_path / "reads.fa.gz"), "wb") as o:
        o.write(f.read())
    out, err = run_cli(CLI, flags_of(run) + ["--batch-pairs", 700, "-qs", pref, "-fa", tmp_path / "reads.fa.gz", "-o", tmp_path / "gz"], tmp_path)
    assert out.decode().splitlines() == exp["kam"]
    assert np.array_equal(np.fromfile(tmp_path / "gz.trkmc.ar", "<u8")[1:], exp["trkmc"])
    r = subprocess.run([CLI, *map(str, flags_of(run) + ["--batch-pairs", 700, "-qs", pref, "-fa", "/dev/stdin", "-o", tmp_path / "pipe"])],
                       input=open(fa, "rb").read(), capture_output=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr.decode()[-1000:]
    assert r.stdout.decode().splitlines() == exp["kam"]
    assert np.array_equal(np.fromfile(tmp_path / "pipe.trkmc.ar", "<u8")[1:], exp["trkmc"])
