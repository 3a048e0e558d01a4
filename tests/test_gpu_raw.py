"""GPU tests of the raw FASTA/FASTQ ingestion (dbt_upload_raw): the device-side record splitter and
neighbour pairing against the Python restatement of the reference's read loop (cpp:1918-1976), and the whole path
(raw text -> records, counts) against the ASCII path on the same pairs.  Bit-exact."""
import random

import numpy as np
import pytest

from common import synthetic_case
from oracle_abi import abi, rpgg_io
from test_host import ref_pairing

pytestmark = pytest.mark.gpu


def fastx_text(seqs_pairs, is_fastq, rng, swap_frac=0.3, single_every=0, short_every=0, suffix=True):
    """seqs_pairs: [(title, first-arriving mate, second-arriving mate)] -> text with the mates as neighbours; optionally
    singletons (a title met once) and short mates in between"""
    lines = []
    mark = "@" if is_fastq else ">"

    def rec(title, seq):
        out = [mark + title, seq]
        if is_fastq:
            out += ["+", "".join(rng.choice("FGH:,#") for _ in seq)]
        return out
    for i, (t, a, b) in enumerate(seqs_pairs):
        if single_every and i % single_every == 1:
            lines += rec(f"lonely{i}/1" if suffix else f"lonely{i}", a[:97])
        if short_every and i % short_every == 2:
            lines += rec(f"short{i}/1", a[:31]) + rec(f"short{i}/2", b)
        s1, s2 = ("/1", "/2") if rng.random() > swap_frac else ("/2", "/1")
        if not suffix:
            s1 = s2 = ""
        lines += rec(t + s1, a) + rec(t + s2, b)
    return ("\n".join(lines) + "\n").encode(), lines


def device_pairs(text, d, po):
    """[(title line, late mate seq, early mate seq)] of the pairs the device formed, from pair_offset"""
    out = []
    for o in po:
        ls = text[int(o):].split(b"\n", 9)
        L = 4 if text[:1] == b"@" else 2
        out.append((ls[L].decode(), ls[L + 1].decode(), ls[1].decode()))
    return out


@pytest.mark.parametrize("is_fastq", [False, True])
@pytest.mark.parametrize("single_every,short_every", [(0, 0), (7, 0), (5, 9)])
def test_raw_split_pairs_like_the_reference(is_fastq, single_every, short_every):
    rng = random.Random(11 + single_every + short_every + is_fastq)
    ix, seqs, _ = synthetic_case(nloci=12, seed=5, npairs=1500, read_seed=9, frac_locus=0.5)
    pairs = [(f"read{i}", seqs[2 * i + 1], seqs[2 * i]) for i in range(len(seqs) // 2)]
    text, lines = fastx_text(pairs, is_fastq, rng, single_every=single_every, short_every=short_every)
    k, cth = 21, 30
    p = abi.default_params(k=k, cth=cth, flags=abi.F_KAM, max_batch_pairs=4096)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    info, po, ph, so, recs, cnt = ctx.run_raw(text, is_fastq, cth + k - 1)
    assert info["flags"] == 0
    exp, db = ref_pairing(lines, is_fastq, cth + k - 1, 10 ** 9, return_db=True)
    # the reference's pairs (title incl. marker, pruned; seq1 = late mate, seq2 = early mate), in its order
    exp_pairs = [(t, s1, s2) for (t, s1, s2, *_q) in exp]
    got = device_pairs(text, info, po)
    prune = lambda t: t[:-2] if len(t) >= 2 and t[-2] == "/" and t[-1] in "12" else t
    assert [(prune(t), a, b) for t, a, b in got] == exp_pairs
    assert info["nrecords"] == len(lines) // (4 if is_fastq else 2)
    assert info["nsingle"] == len(db)
    singles = sorted(prune(text[int(o):].split(b"\n", 1)[0].decode()) for o in so)
    assert singles == sorted(db)
    L = ctx.L
    for (t, _, _), h in zip(got, ph):
        assert int(h) == L.dbt_title_hash(t.encode(), len(t))
    # the whole path: same records and counts as the ASCII path fed with the reference's pairing
    trk, kmc, nmap, _ = ctx.download()
    ctx.close()
    flat = []
    for t, s1, s2 in exp_pairs:
        flat += [s1, s2]
    bases, off = rpgg_io.seqs_to_batch(flat)
    ctx2 = abi.Context(p, abi.make_index_struct(ix))
    recs2, cnt2 = ctx2.run(bases, off)
    trk2, kmc2, nmap2, _ = ctx2.download()
    ctx2.close()
    assert recs.tobytes() == recs2.tobytes()
    assert np.array_equal(cnt, cnt2) and np.array_equal(trk, trk2) and np.array_equal(kmc, kmc2) and np.array_equal(nmap, nmap2)
    assert int(trk.sum()) > 0


def test_raw_split_flags_runs_of_equal_titles_and_bad_input():
    ix, seqs, _ = synthetic_case(nloci=4, seed=5, npairs=10, read_seed=9)
    p = abi.default_params(k=21, cth=30, flags=abi.F_KAM, max_batch_pairs=256)
    ctx = abi.Context(p, abi.make_index_struct(ix))
    text = b"".join(b">x/1\n" + s.encode() + b"\n" for s in seqs[:3])          # three records, one title
    info, *_ = ctx.run_raw(text, False, 50)
    assert info["flags"] & abi.RAW_SLOWPATH
    # text that does not end with a newline is refused
    with pytest.raises(abi.DbtError):
        ctx.run_raw(b">a\nACGT", False, 50)
    # empty text: an empty batch
    info, po, ph, so, recs, cnt = ctx.run_raw(b"", False, 50)
    assert info["npairs"] == 0
    # a read longer than the context was built for: DBT_E_CAPACITY
    long_text = (">a/1\n" + "ACGT" * 60 + "\n>a/2\n" + "ACGT" * 60 + "\n").encode()
    with pytest.raises(abi.DbtError):
        ctx.run_raw(long_text, False, 50)
    # more singletons than the caller's capacity -> overflow flag, nothing uploaded
    many = b"".join(f">s{i}\n".encode() + seqs[0].encode() + b"\n" for i in range(64))
    info, *_ = ctx.run_raw(many, False, 50, single_cap=8)
    assert info["flags"] & abi.RAW_OVERFLOW
    ctx.close()
