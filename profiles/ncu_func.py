import csv,io,os,re,subprocess,sys,tempfile,collections
rep, lib, kern = sys.argv[1:4]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
lines=None
for f in os.listdir(tmp):
    if not f.endswith(".cubin"): continue
    s = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    cur, out = None, {}
    for ln in s.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m: cur = m.group(1); out[cur] = []; continue
        if cur is not None: out[cur].append(ln)
    for name, body in out.items():
        if kern in name: lines = body
seq, curfn, curline = [], "kernel", "?"
for ln in lines:
    m = re.match(r"\s*(\$?_Z\S+):\s*$", ln)
    if m:
        nm=m.group(1)
        mm=re.search(r"_ZN3dbt\d+([a-z_]+)", nm.rsplit("$",1)[-1]) if "$" in nm else None
        curfn = mm.group(1) if mm else "kernel"
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: curline = os.path.basename(m.group(1)) + ":" + m.group(2); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln): seq.append((curfn,curline))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + (["-k", "regex:" + sys.argv[4]] if len(sys.argv) > 4 else []), capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
names=["Instructions Executed","# Samples","stall_no_inst","stall_long_sb","stall_wait","stall_short_sb","stall_branch_resolving"]
cols=[h.index(n) for n in names]
data = [r for r in rows[hi + 1:] if r and r[0].startswith("0x")]
agg=collections.defaultdict(lambda:[0]*len(cols)); st=collections.Counter()
for r,(fn,l) in zip(data,seq):
    st[fn]+=1
    for k,c in enumerate(cols): agg[fn][k]+=int(r[c] or 0)
tot=[sum(a[k] for a in agg.values()) for k in range(len(cols))]
print("function            static   inst%   smp%  no_inst long_sb  wait short_sb branch")
for fn,a in sorted(agg.items(), key=lambda kv:-kv[1][1]):
    print(f"{fn:18s} {st[fn]:6d} {100*a[0]/tot[0]:7.2f} {100*a[1]/tot[1]:6.2f} "+" ".join(f"{100*a[k]/tot[1]:7.2f}" for k in range(2,len(cols))))
print("total inst",tot[0],"samples",tot[1])
