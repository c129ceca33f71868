#!/usr/bin/env python
"""per-iteration warp-instruction counts by source line: python tools/ncu_lines2.py src.csv lib.so kernel n_iterations [top]"""
import csv, re, os, sys, collections, subprocess, tempfile
src_csv, so, kname, nit = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and kname in l][0]
inl = re.compile(r'//## File "([^"]+)", line (\d+)'); ins = re.compile(r'^\s+/\*[0-9a-f]{4,}\*/\s+(.*?);')
lines = []; cur = ("?", 0)
for l in dis[start + 1:]:
    if l.startswith("//---------------------") and ".text." in l: break
    m = inl.search(l)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = ins.match(l)
    if m: lines.append((cur, m.group(1)))
rows = list(csv.reader(open(src_csv)))
hi = [i for i, r in enumerate(rows) if "Instructions Executed" in r][0]
h = rows[hi]; iI = h.index("Instructions Executed"); iW = h.index("Warp Stall Sampling (All Samples)")
agg = collections.Counter(); ldl = collections.Counter(); st = collections.Counter()
for r, (ln, txt) in zip(rows[hi + 1:], lines):
    n = int(r[iI]); agg[ln] += n; st[ln] += int(r[iW])
    if "LDL" in txt or "STL" in txt: ldl[ln] += n
ts = sum(st.values())
print("per-iteration warp instructions by line:")
for ln, n in agg.most_common(top):
    print("%8.1f  %5.1f%% stall  %s:%d   local=%.1f" % (n / nit, 100 * st[ln] / ts, ln[0], ln[1], ldl[ln] / nit))
print("total/iter %.0f  local/iter %.0f" % (sum(agg.values()) / nit, sum(ldl.values()) / nit))
