#!/usr/bin/env python3
"""Golden vectors for `dbt-ktools serialize` (tests/test_serialize.py): a small text RPGG whose k-mers repeat inside a
locus, across two and across three loci (the vv-list path of readKmerIndex, src/kmerIO.hpp:47-78) is serialized by the
compiled reference (oracle/_ref/ktools serialize, src/kmertools.cpp:221-345); the text inputs (gz) and the SHA-256 of
the three binaries are committed.  Run from the repo root: python tests/golden/make_golden_serialize.py"""
import gzip, hashlib, json, os, random, shutil, subprocess, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
OUT = os.path.join(ROOT, "tests", "golden", "serialize")
KTOOLS = os.path.join(ROOT, "oracle", "_ref", "ktools")


def main():
    rng = random.Random(20261105)
    os.makedirs(OUT, exist_ok=True)
    tmp = tempfile.mkdtemp()
    nloci = 5
    shared2 = [rng.getrandbits(42) for _ in range(40)]
    shared3 = [rng.getrandbits(42) for _ in range(15)]

    def locus_kmers(l, n, width):
        ks = [rng.getrandbits(width) for _ in range(n)]
        if l in (0, 1): ks += shared2
        if l in (1, 2, 4): ks += shared3
        ks += ks[:7]                      # duplicate lines inside a locus
        rng.shuffle(ks)
        return ks
    for ext, n, width in (("tr", 300, 42), ("fl", 200, 42), ("tre", 250, 44)):
        with open(os.path.join(tmp, f"pan.{ext}.kmers"), "w") as f:
            for l in range(nloci):
                f.write(f">locus{l}\n")
                for km in locus_kmers(l, n + 13 * l, width):
                    f.write(f"{km}\t0\n")
    r = subprocess.run([KTOOLS, "serialize", os.path.join(tmp, "pan")], capture_output=True)
    assert r.returncode == 0, r.stderr
    meta = {"generator": "oracle/_ref/ktools serialize (reference src/kmertools.cpp:221-345, g++ -O3, libstdc++ of this image)", "sha256": {}}
    for ext in ("kmers.dbi", "fl.kdb", "tre.kdb"):
        meta["sha256"][f"pan.{ext}"] = hashlib.sha256(open(os.path.join(tmp, f"pan.{ext}"), "rb").read()).hexdigest()
    for ext in ("tr", "fl", "tre"):
        with open(os.path.join(tmp, f"pan.{ext}.kmers"), "rb") as f, gzip.GzipFile(os.path.join(OUT, f"pan.{ext}.kmers.gz"), "wb", mtime=0) as o:
            o.write(f.read())
    json.dump(meta, open(os.path.join(OUT, "meta.json"), "w"), indent=1)
    shutil.rmtree(tmp)


if __name__ == "__main__":
    main()
