"""tools/ctl_probe.py -- where the time of one closed-loop MPC period goes (run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")]
import numpy as np
import torch
from qrw_b200 import workloads as wl
from qrw_b200.controller import BatchedController
from qrw_b200.batched import BatchedQPWBC

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
cw = wl.PlannerWorkload(B, 9, change_every=1000, yaw0=0.0, speed=0.5)
cin = [[torch.from_numpy(a).cuda() for a in cw.tick(k)] for k in range(40)]
ctl = BatchedController(B)
L = ctl._L
import ctypes as C
def wbc_info():
    it = np.empty(B, np.int32); st = np.empty(B, np.int32); rho = np.empty(B)
    L.qrw_wbc_info_host(C.c_void_p(ctypes_handle), it.ctypes.data, st.ctypes.data, rho.ctypes.data)
    return it, st
for k in range(40):
    e0, e1, e2, e3 = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    e0.record()
    ctl.planner.run(k, *cin[k])
    e1.record()
    if k % 10 == 0:
        ctl.mpc.run(k, ctl.out["xref"], ctl.out["fsteps"])
    e2.record()
    o = ctl.out
    f_ptr, stride = (ctl.mpc_result.data_ptr() + 96, 384) if k else (ctl.default_forces.data_ptr(), 12)
    L.qrw_wbctick_run_ctrl_device(ctl.wbc._h, ctl.h_ref, C.c_void_p(cin[k][2].data_ptr()), C.c_void_p(f_ptr), stride,
                                  C.c_void_p(o["contacts"].data_ptr()), C.c_void_p(o["p_cmd"].data_ptr()),
                                  C.c_void_p(o["v_cmd"].data_ptr()), C.c_void_p(o["a_cmd"].data_ptr()), None)
    e3.record()
    torch.cuda.synchronize()
    if k >= 10:
        print(f"k={k:3d} planner {e0.elapsed_time(e1):7.3f} ms  mpc {e1.elapsed_time(e2):7.3f} ms  wbc tick {e2.elapsed_time(e3):7.3f} ms")
