#!/usr/bin/env python3
"""Per-CUDA-source-line instruction/stall summary of one kernel from an .ncu-rep.

usage: ncu_lines.py REPORT.ncu-rep LIB.so KERNEL_SUBSTRING [TOPN]
Joins `ncu --page source --csv` (SASS rows, in address order) with `nvdisasm -g` line info of the
same kernel extracted from the shared library (the library must be the build that was profiled).
"""
import csv, io, os, re, subprocess, sys, tempfile

rep, lib, kern = sys.argv[1:4]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
lines = None
for f in os.listdir(tmp):
    if not f.endswith(".cubin"):
        continue
    s = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    # split into functions
    cur, out = None, {}
    for ln in s.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            cur = m.group(1); out[cur] = []; continue
        if cur is not None:
            out[cur].append(ln)
    for name, body in out.items():
        if kern in name:
            lines = body; kname = name
if lines is None:
    sys.exit("kernel not found in " + lib)
# walk: track current (file,line), collect one entry per instruction
seq, curline = [], "?"
for ln in lines:
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        curline = os.path.basename(m.group(1)) + ":" + m.group(2)
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        seq.append(curline)
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + (["-k", "regex:" + os.environ["NCU_K"]] if os.environ.get("NCU_K") else []), capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ci, si, ti = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
data = [r for r in rows[hi + 1:] if len(r) > ci and r[0].startswith("0x")]
if len(data) != len(seq):
    print(f"warning: {len(data)} profiled SASS rows vs {len(seq)} disassembled instructions", file=sys.stderr)
agg = {}
for r, l in zip(data, seq):
    a = agg.setdefault(l, [0, 0, 0])
    a[0] += int(r[ci] or 0); a[1] += int(r[si] or 0); a[2] += int(r[ti] or 0)
tot = sum(a[0] for a in agg.values()); tots = sum(a[1] for a in agg.values())
print(f"kernel {kname}\ntotal warp instructions {tot}, samples {tots}")
src_cache = {}
def src(l):
    f, n = l.rsplit(":", 1)
    for root in (os.path.dirname(os.path.abspath(lib)) + "/csrc", "."):
        p = os.path.join(root, f)
        if os.path.exists(p):
            if p not in src_cache: src_cache[p] = open(p).read().splitlines()
            try: return src_cache[p][int(n) - 1].strip()[:100]
            except Exception: return ""
    return ""
print(f"{'inst%':>6} {'smp%':>6} {'thr/inst':>8}  line")
for l, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    print(f"{100*a[0]/max(tot,1):6.2f} {100*a[1]/max(tots,1):6.2f} {a[2]/max(a[0],1):8.1f}  {l:22s} {src(l)}")
