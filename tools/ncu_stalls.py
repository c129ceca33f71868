"""tools/ncu_stalls.py -- per-instruction stall samples of one kernel from `ncu --page source --csv --print-source sass`.
usage: ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv ; python tools/ncu_stalls.py src.csv [lo_hex hi_hex]
Prints the total samples by stall reason and the instructions with the most samples (optionally inside an address window)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {n: i for i, n in enumerate(hdr)}
body = rows[2:]
base = int(body[0][0], 16)
lo = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 60
reasons = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
tot = {r: 0 for r in reasons}
items = []
nsamp = 0
for r in body:
    off = int(r[0], 16) - base
    if not (lo <= off < hi):
        continue
    s = int(r[ix["# Samples"]] or 0)
    nsamp += s
    for q in reasons:
        tot[q] += int(r[ix[q]] or 0)
    items.append((s, off, r[1].strip(), {q: int(r[ix[q]] or 0) for q in reasons}, int(r[ix["Instructions Executed"]] or 0)))
print("samples", nsamp)
for q, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]:
    print(f"  {q:28s} {v:8d} {100.0 * v / max(nsamp, 1):5.1f} %")
items.sort(key=lambda t: -t[0])
for s, off, txt, d, ne in items[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    top = sorted(d.items(), key=lambda kv: -kv[1])[:2]
    print(f"{off:06x} {s:6d} {ne:9d}  {txt[:60]:60s} " + " ".join(f"{k[6:]}={v}" for k, v in top if v))
