"""tools/ab_probe.py -- A/B of the MPC kernel variants in the A/B test library on the bench workload (4096 trot robots, update
ticks): solves/s per variant, and each variant against variant 2 (iteration counts, results).
usage (GPU box): python tools/ab_probe.py 2 4 [3 ...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")]
import numpy as np, torch
from qrw_b200 import BatchedMPC, _lib, workloads as wl

variants = [int(a) for a in sys.argv[1:]] or [2, 4]
AB = os.environ.get("QRW_LIB", _lib._AB_SO)  # QRW_LIB: any other build of the library (then variant 2 only)
B, T = 4096, 8
w = wl.MpcWorkload(B, int(os.environ.get("AB_SEED", "1")), gaits=("trot",), ticks=T + 2)
dev = torch.device("cuda", 0)
ref = None
for v in variants:
    m = BatchedMPC(B, library=AB)
    m.set_kernel(v)
    x0, res, its = None, [], []
    xs = []
    for t in range(T + 1):
        xr, fs = w.inputs(t, x0)
        xs.append((torch.from_numpy(xr).to(dev), torch.from_numpy(fs).to(dev)))
        m.run(t, *xs[-1])
        torch.cuda.synchronize()
        r = m.get_latest_result()
        res.append(r.copy()); its.append(m.info()["iters"].copy())
        x0 = (ref[0][t] if ref else r)[:, :12].copy()  # every variant replays variant[0]'s history
    if ref is None:
        ref = (res, its)
    # timing: replay ticks 1..T from the state after tick 0
    m2 = BatchedMPC(B, library=AB); m2.set_kernel(v)
    m2.run(0, *xs[0]); torch.cuda.synchronize()
    for t in range(1, 4): m2.run(t, *xs[t])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for t in range(4, T + 1): m2.run(t, *xs[t])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / (T - 3)
    same = [np.mean(a == b) for a, b in zip(its, ref[1])]
    err = max(np.abs(a - b).max() for a, b in zip(res, ref[0]))
    print(f"variant {v}: {ms:.3f} ms per update tick = {B / ms:.1f} k solves/s; vs variant {variants[0]}: same iteration counts {min(same):.4f}, max |diff| {err:.2e}")
