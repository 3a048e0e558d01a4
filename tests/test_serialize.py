"""CPU tests of SURVEY §8(f) row 2: index construction (`dbt-ktools serialize` == the reference's `ktools serialize`,
src/kmertools.cpp:221-345) and the legacy v1.3 containers (PREF.kmerDBi.umap/.vv, PREF.ntr.kmers) the CLI loads.
Golden files under tests/golden/serialize were produced by the compiled reference (make_golden_serialize.py)."""
import gzip
import hashlib
import json
import os
import shutil
import struct
import subprocess

import numpy as np
import pytest

from common import GOLDEN, synth
from oracle_abi import PKG_DIR

KTOOLS = os.path.join(PKG_DIR, "bin", "dbt-ktools")
HOSTCHECK = os.path.join(PKG_DIR, "bin", "dbt-hostcheck")
REF_KTOOLS = os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "ktools")
SER = os.path.join(GOLDEN, "serialize")
pytestmark = pytest.mark.skipif(not os.path.exists(KTOOLS), reason="bin/dbt-ktools not built (run __graft_entry__.build())")


def sha(fn):
    return hashlib.sha256(open(fn, "rb").read()).hexdigest()


def test_serialize_matches_reference_golden(tmp_path):
    """text RPGG of the golden (3 loci, shared k-mers -> vv lists, duplicate lines) -> the three binaries ktools wrote"""
    meta = json.load(open(os.path.join(SER, "meta.json")))
    for ext in ("tr.kmers", "fl.kmers", "tre.kmers"):
        with gzip.open(os.path.join(SER, f"pan.{ext}.gz"), "rb") as f, open(tmp_path / f"pan.{ext}", "wb") as o:
            o.write(f.read())
    r = subprocess.run([KTOOLS, "serialize", str(tmp_path / "pan")], capture_output=True)
    assert r.returncode == 0, r.stderr
    for ext in ("kmers.dbi", "fl.kdb", "tre.kdb"):
        assert sha(tmp_path / f"pan.{ext}") == meta["sha256"][f"pan.{ext}"], ext


def write_text_kmers(fn, per_locus):
    with open(fn, "w") as f:
        for i, ks in enumerate(per_locus):
            f.write(f">locus{i}\n")
            for km in ks:
                f.write(f"{int(km)}\t0\n")


@pytest.mark.skipif(not os.path.exists(REF_KTOOLS), reason="oracle/_ref/ktools not built (needs /root/reference)")
@pytest.mark.parametrize("seed,nloci", [(1, 1), (2, 40), (3, 300)])
def test_serialize_matches_live_reference(tmp_path, seed, nloci):
    """seeded synthetic text RPGGs: our three files are byte-identical to the live reference's"""
    rng = np.random.default_rng(seed)
    loci = synth.make_loci(nloci, seed=seed, nhap=3, flank=120, share_frac=0.5, share_len=90, tr_median=150)
    a, b = tmp_path / "a", tmp_path / "b"
    a.mkdir(); b.mkdir()
    synth.build_rpgg(loci, str(a / "pan"), k=21, fsi=120, fso=120, write_text=True)
    # a .tre.kmers (the generator writes none): random (k+1)-mers, some repeated within and across loci
    tre = []
    pool = rng.integers(0, 1 << 44, size=200, dtype=np.uint64)
    for _ in range(nloci):
        n = int(rng.integers(0, 60))
        tre.append(list(rng.choice(pool, size=n)) + list(rng.integers(0, 1 << 44, size=n // 2, dtype=np.uint64)))
    write_text_kmers(a / "pan.tre.kmers", tre)
    for ext in ("kmers.dbi", "fl.kdb", "tre.kdb"):
        if os.path.exists(a / f"pan.{ext}"):
            os.remove(a / f"pan.{ext}")
    for ext in ("tr.kmers", "fl.kmers", "tre.kmers"):
        shutil.copy(a / f"pan.{ext}", b / f"pan.{ext}")
    r1 = subprocess.run([REF_KTOOLS, "serialize", str(a / "pan")], capture_output=True)
    assert r1.returncode == 0, r1.stderr
    r2 = subprocess.run([KTOOLS, "serialize", str(b / "pan")], capture_output=True)
    assert r2.returncode == 0, r2.stderr
    for ext in ("kmers.dbi", "fl.kdb", "tre.kdb"):
        assert open(a / f"pan.{ext}", "rb").read() == open(b / f"pan.{ext}", "rb").read(), ext


def test_serialize_errors(tmp_path):
    assert subprocess.run([KTOOLS, "serialize", str(tmp_path / "missing")], capture_output=True).returncode == 1
    (tmp_path / "pan.tr.kmers").write_text("123\t0\n>l0\n")       # k-mer line before the first header: the reference indexes [-1]
    (tmp_path / "pan.fl.kmers").write_text(">l0\n")
    (tmp_path / "pan.tre.kmers").write_text(">l0\n")
    assert subprocess.run([KTOOLS, "serialize", str(tmp_path / "pan")], capture_output=True).returncode == 1


def test_legacy_v13_containers_load(tmp_path):
    """PREF.kmerDBi.umap / .vv / .ntr.kmers (v1.3 release layout) load as the same index the new files give.
    Fixture files of that layout ship in the reference's test/QC/input; their byte layout (8 + n*16, 8 + 4n) is what
    read_legacy_dbi parses -- here the same layout is written from the golden index and must load identically."""
    d = os.path.join(GOLDEN, "multi")
    for ext in ("tr.kmers", "kmers.dbi", "fl.kdb"):
        with gzip.open(os.path.join(d, f"pan.{ext}.gz"), "rb") as f, open(tmp_path / f"pan.{ext}", "wb") as o:
            o.write(f.read())
    new = subprocess.run([HOSTCHECK, "index-sum", str(tmp_path / "pan")], capture_output=True)
    assert new.returncode == 0, new.stderr
    raw = open(tmp_path / "pan.kmers.dbi", "rb").read()
    nk = struct.unpack_from("<Q", raw, 0)[0]
    keys = np.frombuffer(raw, "<u8", nk, 8)
    vals = np.frombuffer(raw, "<u4", nk, 8 + 8 * nk)
    nvv = struct.unpack_from("<Q", raw, 8 + 12 * nk)[0]
    vv = np.frombuffer(raw, "<u4", nvv, 16 + 12 * nk)
    leg = tmp_path / "leg"
    leg.mkdir()
    shutil.copy(tmp_path / "pan.tr.kmers", leg / "pan.tr.kmers")
    kv = np.empty((nk, 2), "<u8"); kv[:, 0] = keys; kv[:, 1] = vals
    with open(leg / "pan.kmerDBi.umap", "wb") as f:
        f.write(struct.pack("<Q", nk)); f.write(kv.tobytes())
    with open(leg / "pan.kmerDBi.vv", "wb") as f:
        f.write(struct.pack("<Q", nvv)); f.write(vv.tobytes())
    # flank k-mers as v1.3 text (PREF.ntr.kmers), from fl.kdb
    rawf = open(tmp_path / "pan.fl.kdb", "rb").read()
    nl = struct.unpack_from("<Q", rawf, 0)[0]
    fn = np.frombuffer(rawf, "<u8", nl, 8)
    fk = np.frombuffer(rawf, "<u8", int(fn.sum()), 16 + 8 * nl)
    per, o = [], 0
    for n in fn:
        per.append(fk[o:o + int(n)]); o += int(n)
    write_text_kmers(leg / "pan.ntr.kmers", per)
    old = subprocess.run([HOSTCHECK, "index-sum", str(leg / "pan")], capture_output=True)
    assert old.returncode == 0, old.stderr
    assert old.stdout == new.stdout and b"nloci" in new.stdout
    # truncated legacy files fail loudly
    os.truncate(leg / "pan.kmerDBi.umap", 8 + 16 * nk - 3)
    assert subprocess.run([HOSTCHECK, "index-sum", str(leg / "pan")], capture_output=True).returncode == 1


REF_QC = "/root/reference/test/QC/input"


@pytest.mark.skipif(not os.path.exists(REF_QC), reason="reference fixtures not present on this box")
def test_legacy_fixture_of_the_reference_loads(tmp_path):
    """the v1.3 fixture in the reference's own test tree (pan.kmerDBi.umap 23,992 B = 8 + 1499*16; .vv n = 0): every key
    of pan.tr.kmers / pan.ntr.kmers is found with value locus<<1 (1 locus)"""
    for fn in ("pan.kmerDBi.umap", "pan.kmerDBi.vv", "pan.tr.kmers", "pan.ntr.kmers"):
        shutil.copy(os.path.join(REF_QC, fn), tmp_path / fn)
    r = subprocess.run([HOSTCHECK, "index-sum", str(tmp_path / "pan")], capture_output=True)
    assert r.returncode == 0, r.stderr
    f = dict(zip(r.stdout.split()[::2], r.stdout.split()[1::2]))
    text = set()
    for fn in ("pan.tr.kmers", "pan.ntr.kmers"):
        text |= {int(l.split()[0]) for l in open(tmp_path / fn) if not l.startswith(">")}
    assert int(f[b"nk_dbi"]) == len(text) == 1499 and int(f[b"nvv"]) == 0 and int(f[b"nloci"]) == 1
    assert int(f[b"keysum"]) == sum(text) % (1 << 64)
