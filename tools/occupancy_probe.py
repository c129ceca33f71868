"""tools/occupancy_probe.py -- MPC kernel: time per CTA-iteration with one and with two CTAs per SM (GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")]
import numpy as np, torch
from qrw_b200 import BatchedMPC, workloads as wl

for B in (74, 148, 296, 592, 1184, 4096):
    w = wl.MpcWorkload(B, 1, gaits=("trot",), ticks=4)
    m = BatchedMPC(B)
    xr, fs = w.inputs(0)
    m.run(0, xr, fs)
    x0 = m.get_latest_result()[:, :12].copy()
    for t in (1, 2):
        xr, fs = w.inputs(t, x0)
        dx, df = torch.from_numpy(xr).cuda(), torch.from_numpy(fs).cuda()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(); m.run(t, dx, df); e1.record(); torch.cuda.synchronize()
        it = m.info()["iters"]
        x0 = m.get_latest_result()[:, :12].copy()
    ms = e0.elapsed_time(e1)
    ncta = min(B, 296)
    # lower bound of the launch: the longest robot alone; average: all CTA-iterations spread over the CTAs
    print(f"B={B:5d} CTAs={ncta:3d} ms={ms:8.3f} sum_iters={it.sum():8d} max_iters={it.max():5d}  us per CTA-iteration (avg) = {1e3*ms*ncta/it.sum():6.3f}"
          f"   us per iteration of the longest robot <= {1e3*ms/it.max():6.3f}")
    m.close()
