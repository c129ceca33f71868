#!/usr/bin/env python
"""tools/make_golden.py -- regenerate tests/golden/*.npz (small input/output fixtures of the hot path).

The reference cannot be built or imported here (OSQP, Eigen, Boost.Python, Pinocchio are absent: SURVEY.md F4),
so the fixtures are produced by the CPU oracle (oracle/, the restatement of src/MPC.cpp, src/QPWBC.cpp,
src/InvKin.cpp + OSQP v0.6) from the seeded workload generators.  They pin (a) the oracle against silent
drift and (b) the CUDA path on the GPU box, where neither /root/reference nor a rebuild of these inputs is
needed.  "Parity unpinned" w.r.t. an OSQP binary still applies.

    python tools/make_golden.py          # rewrites tests/golden/
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")):
    sys.path.insert(0, p)

from oracle import bindings as ob  # noqa: E402
from qrw_b200 import workloads as wl  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def mpc_case(name, B, seed, gaits, ticks, single=None):
    o = ob.OracleMPC(B)
    rec = {}
    x0 = None
    if single is None:
        w = wl.MpcWorkload(B, seed, gaits=gaits, ticks=ticks + 1)
    for t in range(ticks):
        xr, fs = single if single is not None else w.inputs(t, x0)
        o.run(t, xr, fs, 8)
        r, info = o.get_latest_result(), o.info()
        rec[f"xref{t}"], rec[f"fsteps{t}"], rec[f"result{t}"] = xr, fs, r
        rec[f"gait{t}"], rec[f"sgait{t}"] = o.get_gait(), o.get_Sgait()
        rec[f"iters{t}"], rec[f"status{t}"], rec[f"rho{t}"] = info["iters"], info["status"], info["rho"]
        x0 = r[:, :12].copy()
    rec["ticks"] = np.array(ticks)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(name, "iters", [rec[f"iters{t}"].tolist()[:4] for t in range(ticks)])


def main():
    os.makedirs(OUT, exist_ok=True)
    mpc_case("mpc_config1_single_trot", 1, 0, None, 2, single=wl.config1_single_trot())
    mpc_case("mpc_trot_b8", 8, 1, ("trot",), 3)
    mpc_case("mpc_mixed_b10", 10, 2, wl.GAITS, 3)
    # QPWBC: three consecutive calls (warm start, rho carried over)
    B = 16
    o = ob.OracleWBC(B)
    rec = {}
    for c in range(3):
        M, Jc, f, R = wl.wbc_inputs(B, seed=30 + c)
        o.run(M, Jc, f, R, 4)
        fr, dd, H = o.results()
        rec.update({f"M{c}": M, f"Jc{c}": Jc, f"f_cmd{c}": f, f"RNEA{c}": R, f"f_res{c}": fr, f"ddq_res{c}": dd, f"H{c}": H,
                    f"iters{c}": o.info()["iters"]})
    np.savez_compressed(os.path.join(OUT, "wbc_b16.npz"), **rec)
    # InvKin
    args = wl.invkin_inputs(B, seed=4)
    outs = ob.invkin(*args)
    names = ["contacts", "goals", "vgoals", "agoals", "posf", "vf", "wf", "af", "Jf"]
    np.savez_compressed(os.path.join(OUT, "invkin_b16.npz"), **dict(zip(names, args)),
                        ddq=outs[0], dq_cmd=outs[1], q_step=outs[2])
    # whole-body tick (Solo12 kinematics / RNEA + InvKin + QPWBC + tau_ff), two calls
    o = ob.OracleWBC(B)
    rec = {}
    names = ["q", "dq", "f_cmd", "contacts", "pgoals", "vgoals", "agoals"]
    for c in range(2):
        args = wl.wbc_tick_inputs(B, seed=50 + c, level_base=(c == 0))
        r = ob.wbc_tick(o, *args, nthreads=4)
        rec.update({f"{n}{c}": a for n, a in zip(names, args)})
        rec.update({f"{k}{c}": v for k, v in r.items()})
    np.savez_compressed(os.path.join(OUT, "wbc_tick_b16.npz"), **rec)
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
