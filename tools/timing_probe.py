"""tools/timing_probe.py -- cycles per segment of an ADMM iteration of mpc_v2_kernel (probe build -DQRW_V2_TIMING).
Build the probe library first:  nvcc <flags> -DQRW_V2_TIMING -o /tmp/libqrw_timing.so csrc/qrw_b200.cu ; then on the GPU box
QRW_LIB=/path/libqrw_timing.so python tools/timing_probe.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")]
import numpy as np
from qrw_b200 import _lib
_lib._SO = os.environ.get("QRW_LIB", _lib._SO)
from qrw_b200 import BatchedMPC, workloads as wl

for B in (148, 296, 4096):
    w = wl.MpcWorkload(B, 1, gaits=("trot",), ticks=4)
    m = BatchedMPC(B)
    xr, fs = w.inputs(0)
    m.run(0, xr, fs)
    x0 = m.get_latest_result()[:, :12].copy()
    m.enable_debug()
    xr, fs = w.inputs(1, x0)
    m.run(1, xr, fs)
    d = m.debug()["bounds"]  # last 192 entries of the dump; the probe build parks its counters at [183:189]
    ph = d[:, 171:178] / d[:, 188][:, None]
    print(f"B={B:5d}: cycles per iteration by phase (1)..(7): " + " ".join(f"{x:6.0f}" for x in ph.mean(0)))
    cs = d[:, 178:182] / d[:, 188][:, None]
    print(f"B={B:5d}: cycles per iteration inside the chain: elimination {cs[:,0].mean():6.0f} | middle {cs[:,1].mean():6.0f} | v products {cs[:,2].mean():6.0f} | back-substitution {cs[:,3].mean():6.0f}")
    t = d[:, 183:192]
    it = t[:, 5]
    per = t[:, :5] / it[:, None]
    print(f"B={B:5d}: cycles per SOLVE (mean)  setup before the loop {t[:,6].mean():9.0f} | factorisations {t[:,7].mean():9.0f} | residual checks {t[:,8].mean():9.0f}"
          f" | iterations {(t[:,0]+t[:,3]+t[:,4]).mean():10.0f}")
    print(f"B={B:5d}: cycles per iteration (mean over robots)  phases 1-4 {per[:,0].mean():7.0f} | chain {per[:,4].mean():7.0f} | phases 5-7 {per[:,3].mean():7.0f}"
          f" | sum {(per[:,0]+per[:,3]+per[:,4]).mean():7.0f}   (iters mean {it.mean():.0f})")
    m.close()
