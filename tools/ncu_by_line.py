"""tools/ncu_by_line.py -- stall samples of one kernel per SOURCE LINE: joins `ncu --page source --csv --print-source sass`
with `nvdisasm --print-line-info-inline` of the same cubin (instruction offsets).  An instruction inlined from a helper is
attributed to the line of <source file> that (transitively) called it.
usage: python tools/ncu_by_line.py src.csv all_inl.sass <kernel substring> <source file> [top N] [lo:hi:name ...]
The optional ranges print the share of the samples between two source lines (phases of a kernel)."""
import collections
import csv
import re
import sys

csvf, sassf, kern, srcname = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 60
ranges = [a.split(":") for a in sys.argv[6:]]
want = srcname.split('/')[-1]
# offset -> (file, line) inside the kernel's .text section; inlined frames: keep the innermost line of `srcname` if present
off2line, infn, cur, chain = {}, False, None, []
for l in open(sassf):
    if l.startswith('//--------------------- .text.'):
        infn = kern in l
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        chain.append((m.group(1).split('/')[-1], int(m.group(2))))
        continue
    if chain:
        # innermost first; take the OUTERMOST frame that lies in the wanted file, else the innermost frame
        inw = [c for c in chain if c[0] == want]
        cur = inw[-1] if inw else chain[0]
        chain = []
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', l)
    if m:
        off2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(csvf)))
hdr = rows[1]
ix = {n: i for i, n in enumerate(hdr)}
body = rows[2:]
base = int(body[0][0], 16)
reasons = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
agg = collections.defaultdict(lambda: collections.Counter())
tot = 0
for r in body:
    off = int(r[0], 16) - base
    s = int(r[ix["# Samples"]] or 0)
    tot += s
    key = off2line.get(off, ("?", 0))
    agg[key]["samples"] += s
    agg[key]["inst"] += int(r[ix["Instructions Executed"]] or 0)
    for q in reasons:
        agg[key][q] += int(r[ix[q]] or 0)
try:
    src = open(srcname).read().split('\n')
except Exception:
    src = []
print("total samples", tot)
items = sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:top]
for (f, ln), c in sorted(items, key=lambda kv: (kv[0][0], kv[0][1])):
    rs = sorted(((q[6:], c[q]) for q in reasons), key=lambda t: -t[1])[:3]
    text = src[ln - 1].strip()[:70] if src and f == srcname.split('/')[-1] and 0 < ln <= len(src) else ""
    print(f"{f:>18s}:{ln:5d} {c['samples']:8d} {100.0*c['samples']/tot:5.1f}%  " + " ".join(f"{a}={b}" for a, b in rs if b) + "  | " + text)

if ranges:
    print("-- ranges (share of all samples)")
    for lo, hi, name in ranges:
        lo, hi = int(lo), int(hi)
        c = collections.Counter()
        for (f, ln), v in agg.items():
            if f == want and lo <= ln <= hi:
                c.update(v)
        rs = sorted(((q[6:], c[q]) for q in reasons), key=lambda t: -t[1])[:4]
        print(f"{name:>14s} [{lo:4d},{hi:4d}] {c['samples']:8d} {100.0*c['samples']/tot:5.1f}%  " + " ".join(f"{a}={b}" for a, b in rs if b))
