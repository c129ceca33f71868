"""tools/ctl_probe2.py -- closed-loop periods: tick-by-tick vs CUDA-graph replay, MPC iteration counts (GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")]
import numpy as np
import torch
from qrw_b200 import workloads as wl
from qrw_b200.controller import BatchedController

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
P = 10
cw = wl.PlannerWorkload(B, 9, change_every=1000, yaw0=0.0, speed=0.5)
cin = [[torch.from_numpy(a).cuda() for a in cw.tick(k)] for k in range(10 * P)]
for mode in ("ticks", "graph"):
    ctl = BatchedController(B)
    for k in range(10):
        ctl.tick(k, *cin[k])
    if mode == "graph":
        stage = ctl.capture_period()
    for p in range(1, P):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        if mode == "ticks":
            for i in range(10):
                ctl.tick(10 * p + i, *cin[10 * p + i])
        else:
            for i in range(10):
                for dst, src in zip(stage, cin[10 * p + i]):
                    dst[i].copy_(src)
            ctl.replay_period()
        e1.record()
        torch.cuda.synchronize()
        it = ctl.mpc.info()["iters"]
        print(f"{mode} period {p}: {e0.elapsed_time(e1):8.3f} ms   MPC iters mean {it.mean():7.1f} p99 {np.percentile(it, 99):7.1f} max {it.max()}")
    ctl.close()
