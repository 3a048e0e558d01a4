"""Development aid: which mates does the fused / fast threading hand on to thread_kernel, by what happened to them in the reference?
Runs the bench's error model at small scale through the oracle (CIGAR per mate -> category) and through a library built with
-DDBT_DEBUG_DIRTY (tools/build_debug_dirty.sh), whose aln records carry the hand-over mask, and cross-tabulates.  GPU needed."""
import collections
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_abi import Oracle, abi, rpgg_io  # noqa: E402  (development script: the oracle is the categoriser here, nothing is shipped)
import thread_common as tc  # noqa: E402

abi.load_library(os.path.join(ROOT, "danbing-tk_b200", "build", "var", "lib_dirty.so"))
cfg = dict(seed=45, rseed=5, sub=0.002, ins=0.00005, dele=0.00005, npairs=int(sys.argv[1]) if len(sys.argv) > 1 else 40000, nloci=40)
_, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.0, **cfg)
bases, off = rpgg_io.seqs_to_batch(seqs)
P = len(seqs) // 2
pt = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=P + 64, thread=85, maxnc=3)
orecs, *_ = Oracle(ix).run(pt, bases, off)
p = tc.thread_params(aln=1, max_batch_pairs=P + 64, thread=85, maxnc=3)
ctx = abi.Context(p, abi.make_index_struct(ix))
recs, cnt = ctx.run(bases, off)
ctx.close()
dirty = {int(r["h"]["pair_index"]): int(r["h"]["pad"][0]) for r in recs}


def category(a):
    nes, al = int(a["h"]["nes"]), int(a["h"]["alned"])
    t = a["es"][:nes] & 7
    xs = [i for i, v in enumerate(t) if v == 2]
    indel = bool(((t == 3) | (t == 4)).any())
    stars = int((t == 0).sum())
    lead = 0
    while lead < nes and t[lead] == 0:
        lead += 1
    trail = 0
    while trail < nes and t[nes - 1 - trail] == 0:
        trail += 1
    if al == 1 and not xs and not indel and stars == 0:
        return "clean"
    k = [f"alned{al}"]
    if indel:
        k.append("indel")
    elif len(xs) == 1:
        q = xs[0]
        k.append("1X@start" if q < 21 else ("1X@end" if q >= nes - 22 else "1X interior"))
    elif len(xs) >= 2:
        k.append(f"{len(xs)}X " + ("close" if min(b - a for a, b in zip(xs, xs[1:])) <= 21 else "far"))
    else:
        k.append("no edit")
    if stars:
        k.append(f"lead*{lead} trail*{trail} inner*{stars - lead - trail}")
    return " ".join(k)


tab = collections.defaultdict(lambda: [0, 0])
n = nd = 0
for r in orecs:
    pi = int(r["h"]["pair_index"])
    if pi not in dirty:
        continue
    d = dirty[pi]
    for m in range(2):
        c = category(r["aln"][m])
        tab[c][0] += 1
        tab[c][1] += (d >> m) & 1
        n += 1
        nd += (d >> m) & 1
print(f"{n} threaded mates, {nd} handed to thread_kernel ({100 * nd / n:.1f} %)")
for c, (a, b) in sorted(tab.items(), key=lambda x: -x[1][1])[:40]:
    print(f"{b:7d} of {a:7d}  {c}")
