#!/usr/bin/env python
"""tools/ncu_lines.py -- aggregate an ncu source-page CSV (SASS level) by CUDA source line.

usage: ncu -i prof.ncu-rep --page source --csv > src.csv
       python tools/ncu_lines.py src.csv libqrw_b200.so <kernel-name-substring> [top]
Line info comes from `nvdisasm -g` of the cubin inside the .so (compiled with -lineinfo); the two listings
are matched instruction by instruction in order.
"""
import csv, re, subprocess, sys, tempfile, os, collections

src_csv, so, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
# locate the kernel's text section
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and kname in l][0]
lines = []  # (file, line) per instruction, in order
cur = ("?", 0)
inl = re.compile(r'//## File "([^"]+)", line (\d+)')
ins = re.compile(r'^\s+/\*[0-9a-f]{4,}\*/\s+(.*?);')
for l in dis[start + 1:]:
    if l.startswith("//---------------------") and ".text." in l:
        break
    m = inl.search(l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if ins.match(l):
        lines.append(cur)
rows = list(csv.reader(open(src_csv)))
hi = [i for i, r in enumerate(rows) if "Instructions Executed" in r][0]
h = rows[hi]
iI, iW, iS = h.index("Instructions Executed"), h.index("Warp Stall Sampling (All Samples)"), h.index("Source")
data = rows[hi + 1:]
if len(data) != len(lines):
    print(f"warning: {len(data)} profiled instructions vs {len(lines)} disassembled", file=sys.stderr)
agg = collections.defaultdict(lambda: [0, 0, 0])
opagg = collections.defaultdict(lambda: [0, 0])
for r, ln in zip(data, lines):
    try:
        n, s = int(r[iI]), int(r[iW])
    except ValueError:
        continue
    a = agg[ln]
    a[0] += n
    a[1] += s
    a[2] += 1
    op = r[iS].split()[0] if not r[iS].startswith("@") else r[iS].split()[1]
    op = op.split(".")[0]
    opagg[op][0] += n
    opagg[op][1] += s
ti = sum(a[0] for a in agg.values())
ts = sum(a[1] for a in agg.values())
print(f"total warp instructions {ti}, stall samples {ts}")
print("--- by source line (sorted by stall samples)")
for ln, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{100*a[0]/ti:6.2f}% inst {100*a[1]/ts:6.2f}% samples  {ln[0]}:{ln[1]}  ({a[2]} sass)")
print("--- by opcode")
for op, a in sorted(opagg.items(), key=lambda kv: -kv[1][0])[:25]:
    print(f"{100*a[0]/ti:6.2f}% inst {100*a[1]/ts:6.2f}% samples  {op}")
