"""tools/iters_dump.py -- ADMM iteration counts per robot and tick of the bench workload (config 2: 4096 trot robots, seed 1),
for the list-scheduling model of the launch tail (tools/tail_model.py).  usage (GPU box): python tools/iters_dump.py out.npz [ticks]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")]
import numpy as np
from qrw_b200 import BatchedMPC, workloads as wl

out = sys.argv[1]
T = int(sys.argv[2]) if len(sys.argv) > 2 else 12
B = 4096
w = wl.MpcWorkload(B, 1, gaits=("trot",), ticks=T + 2)
m = BatchedMPC(B)
x0, its = None, []
for t in range(T + 1):
    xr, fs = w.inputs(t, x0)
    m.run(t, xr, fs)
    r = m.get_latest_result()
    its.append(m.info()["iters"].copy())
    x0 = r[:, :12].copy()
its = np.array(its)
np.savez_compressed(out, iters=its.astype(np.int16), speed=np.linalg.norm(w.vref[:, :2], axis=1))
print("mean iterations per tick", its.mean(1).round(1))
