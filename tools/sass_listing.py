"""tools/sass_listing.py -- SASS of the shipped MPC kernel's ADMM loop body and of its noinline chain functions, with the
source line of mpc_v2_core.cuh that (transitively) contains each instruction (profiles/r02_sass_mpc_v2_loop_and_chain.txt).
usage: python tools/sass_listing.py [out file]      (needs cuobjdump / nvdisasm and the built library; no GPU)"""
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.environ.get("QRW_LIB", os.path.join(ROOT, "quadruped-reactive-walking_b200", "libqrw_b200.so"))
SRC = os.path.join(ROOT, "quadruped-reactive-walking_b200", "csrc", "mpc_v2_core.cuh")
out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_mpc_v2_loop_and_chain.txt")

with tempfile.TemporaryDirectory() as d:
    subprocess.check_call(["cuobjdump", "-xelf", "all", SO], cwd=d, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    sass = subprocess.run(["nvdisasm", "--print-line-info-inline", cubin], cwd=d, capture_output=True, text=True).stdout.split("\n")

src = open(SRC).read().split("\n")
lo = next(i for i, l in enumerate(src) if "for (iter = 1; iter <= C.max_iter; ++iter)" in l) + 1
hi = next(i for i, l in enumerate(src) if "if (--until_check == 0) {" in l) + 1
c0 = next(i for i, l in enumerate(src) if l.startswith("QRW_HDN void state_elim")) + 1
c1 = next(i for i, l in enumerate(src) if "One full MPC::run for one instance" in l)
fl = next(i for i, l in enumerate(src) if "factor(C, S, R, Rh, S.red[120]" in l) + 1  # the (inlined, rarely taken) factorisation: left out
out = ["# SASS of mpc_v2_kernel (the shipped MPC kernel, v2.11), sm_100a, nvcc 12.9 -O3 -lineinfo:",
       f"#  (1) the noinline chain functions state_elim / state_helper / state_back (mpc_v2_core.cuh lines {c0}..{c1}),",
       f"#  (2) the ADMM loop body, lines {lo}..{hi} (phases 1-4, park / chain call / unpark of the chain warps, phases 5-7) without the",
       f"#      inlined factor() of line {fl} (taken 2-3 times per solve).",
       "# In front of each instruction: the line of mpc_v2_core.cuh that (transitively) contains it (helpers such as lds12, dot12a,",
       "# tmem_ld12 are inlined).  tools/sass_listing.py: cuobjdump -xelf all ; nvdisasm --print-line-info-inline"]
infn, chain, cur, cnt = False, [], None, 0
for l in sass:
    if l.startswith("//--------------------- .text."):
        infn = "mpc_v2_kernel" in l
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        chain.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    if chain:
        inw = [c for c in chain if c[0] == "mpc_v2_core.cuh"]
        cur = inw[-1] if inw else None
        chain = []
    m2 = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", l)
    if m2 and cur and cur[1] != fl and ((lo <= cur[1] <= hi) or (c0 <= cur[1] <= c1)):
        out.append(f"{cur[1]:<5d} | /*{m2.group(1)}*/ {m2.group(2)}")
        cnt += 1
counts = {k: sum(1 for o in out[6:] if re.search(k, o)) for k in ("LDTM", "DFMA", "DADD", "DMUL", r"\bLDS", r"\bSTS", "BAR", r"LDL|STL")}
out.insert(6, "# instruction counts in this listing: " + ", ".join(f"{k} {v}" for k, v in counts.items()) + f" (of {cnt})")
open(out_path, "w").write("\n".join(out) + "\n")
print(cnt, "instructions ->", out_path, counts)
