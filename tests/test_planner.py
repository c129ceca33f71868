"""Planners: Gait, StatePlanner, FootstepPlanner, FootTrajectoryGenerator, feet commands.

CPU: (i) the oracle restatement (oracle/planner_oracle.cpp) pinned by properties the reference's algorithm guarantees
(the reference has no test or golden vector for these classes: PARITY UNPINNED), (ii) the product's planner core
(csrc/planner_core.cuh) compiled by g++ in emulation mode against the oracle on a seeded closed-loop tick sequence.
GPU: the CUDA kernel through the C ABI against the oracle on the same sequence, the device-pointer path, ragged and
full-size batches.  Tolerance for doubles: 1e-9 abs + 1e-9 rel (sin/cos/atan2 and FMA contraction differ between
glibc/g++ and CUDA; the swing polynomials amplify round-off by cancellation); 0/1 gait matrices and flags bit-exact.
"""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import bindings as ob
from qrw_b200 import workloads as wl

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
H_REF = 0.2447495
FIELDS = ("cur_fs", "o_target", "fpos", "fvel", "facc", "p_cmd", "v_cmd", "a_cmd", "t0s", "t_swing", "contacts")


def _p(a):
    return a.ctypes.data_as(_dp)


def _close(a, b, tol=1e-9):
    return np.abs(a - b).max() <= tol * (1.0 + np.abs(b).max())


# ------------------------------------------------------------------------------------------ oracle pins (CPU)
def _trot_rows():
    g = np.zeros((20, 4))
    g[:8] = [1, 0, 0, 1]
    g[8:16] = [0, 1, 1, 0]
    return g


def _idle_inputs(B):
    q = np.zeros((B, 7))
    q[:, 2], q[:, 6] = H_REF, 1.0
    return q, np.zeros((B, 6)), np.zeros((B, 6))


def test_gait_initial_trot_and_period():
    """Gait::initialize -> create_trot + create_gait_f (Gait.cpp:21-36,57-69,111-140); rollGait has period T_gait/dt."""
    P = ob.OraclePlanner(1)
    q, v, r = _idle_inputs(1)
    g0 = P.outputs()["gait"].reshape(4, 20).T
    assert (g0 == _trot_rows()).all()
    assert (P.state()["desired"].reshape(4, 20).T == _trot_rows()).all()  # 16 rows consumed = one full period
    seen = []
    for k in range(0, 160, 10):
        P.run(k, q, v, r, stages=ob.OraclePlanner.GAIT)
        g = P.outputs()["gait"].reshape(4, 20).T
        assert (g[16:] == 0).all() and (g[:16].sum(1) == 2).all()
        seen.append(P.state()["flags"][0, 0])
    assert (g == g0).all()  # 16 rolls later the horizon is back where it started
    assert sum(seen) == 2  # two phase changes per trot period (newPhase, Gait.cpp:232-240)
    past = P.state()["past"].reshape(4, 20).T
    assert (past[:16] == g0[:16][::-1][np.r_[0:16]]).all() or past[:17].sum() == 32


def test_gait_only_rolls_every_k_mpc():
    P = ob.OraclePlanner(1)
    q, v, r = _idle_inputs(1)
    P.run(0, q, v, r, stages=1)
    g = P.outputs()["gait"].copy()
    for k in range(1, 10):
        P.run(k, q, v, r, stages=1)
        assert (P.outputs()["gait"] == g).all()
    P.run(10, q, v, r, stages=1)
    assert not (P.outputs()["gait"] == g).all()


def test_gait_too_small_matrix_is_rejected():
    with pytest.raises(ValueError):  # Gait.cpp:30-31 throws std::invalid_argument
        ob.OraclePlanner(1, N_gait=12)


def test_phase_duration_trot():
    """A trot stance / swing phase lasts T_gait / 2 whatever the row it is asked from (Gait.cpp:142-187)."""
    P = ob.OraclePlanner(1)
    q, v, r = _idle_inputs(1)
    for k in range(0, 400, 10):
        P.run(k, q, v, r, stages=1)
    g = P.outputs()["gait"].reshape(4, 20).T
    for i in range(0, 12):
        for j in range(4):
            t, rem = P.phase_duration(0, i, j, g[i, j])
            assert abs(t - 0.16) < 1e-15 and 1 <= rem <= 8


def test_state_planner_matches_independent_restatement():
    """StatePlanner::computeReferenceStates against the vectorised numpy restatement used for the MPC workloads."""
    B = 64
    W = wl.PlannerWorkload(B, seed=11)
    q7, hv, vr, _ = W.tick(0)
    P = ob.OraclePlanner(B, h_ref=wl.H_REF)
    P.run(0, q7, hv, vr, stages=ob.OraclePlanner.STATE)
    xr = P.outputs()["xref"].reshape(B, 17, 12).transpose(0, 2, 1)
    x0 = np.zeros((B, 12))
    x0[:, 2], x0[:, 6:] = q7[:, 2], hv
    ref = wl.reference_states(x0, vr)
    assert np.abs(xr[:, :, 1:] - ref[:, :, 1:]).max() < 1e-14
    assert np.abs(xr[:, 6:, 0] - hv).max() == 0 and (xr[:, [0, 1, 5], 0] == 0).all() and (xr[:, 2, 0] == q7[:, 2]).all()
    # roll / pitch of column 0 come back from the quaternion (matrixToRpy o toRotationMatrix)
    rng = np.random.default_rng(11)
    _ = rng.uniform(-3, 3, B), rng.uniform(-0.5, 1.5, B), rng.uniform(-0.5, 0.5, B), rng.uniform(-0.7, 0.7, B), rng.random(B)
    roll, pitch = rng.uniform(-0.1, 0.1, B), rng.uniform(-0.1, 0.1, B)
    assert np.abs(xr[:, 3, 0] - roll).max() < 1e-14 and np.abs(xr[:, 4, 0] - pitch).max() < 1e-14


def test_footsteps_at_rest_are_the_shoulders_and_swing_entries_are_exact_zeros():
    P = ob.OraclePlanner(1)
    q, v, r = _idle_inputs(1)
    sh = ob.SHOULDERS34.reshape(4, 3)
    for k in range(0, 200):
        P.run(k, q, v, r, stages=3)
        o = P.outputs()
        g = o["gait"].reshape(4, 20).T
        fs = o["fsteps"].reshape(12, 20).T.reshape(20, 4, 3)
        assert ((fs[:, :, 0] != 0.0) == (g == 1.0)).all()  # what MPC::construct_gait reads back (MPC.cpp:686-701)
        assert np.abs(fs[g == 1.0] - sh[np.nonzero(g == 1.0)[1]]).max() < 1e-15
        assert np.abs(o["o_target"].reshape(4, 3) - sh).max() < 1e-15


def test_footstep_heuristic_symmetry_term():
    """Walking straight at v: a foot lands t_stance/2 * v ahead of its shoulder (FootstepPlanner.cpp:160-163), shifted by
    the distance the base covers until touch-down (:103-121)."""
    P = ob.OraclePlanner(1)
    q, v, r = _idle_inputs(1)
    v[0, 0] = r[0, 0] = 0.4
    P.run(0, q, v, r, stages=3)
    fs = P.outputs()["fsteps"].reshape(12, 20).T.reshape(20, 4, 3)
    # tick 0 rolls the gait once (Gait.cpp:195-196), so row 7 is the first row of the (FR, HL) stance; the base has
    # moved 0.4 * (10*0.002 + 6*0.02) by the end of row 6
    exp_x = ob.SHOULDERS34[3] + 0.16 * 0.5 * 0.4 + 0.4 * (10 * 0.002 + 6 * 0.02)
    assert fs[6, 1, 0] == 0.0 and abs(fs[7, 1, 0] - exp_x) < 1e-15 and abs(fs[7, 1, 1] - ob.SHOULDERS34[4]) < 1e-15
    assert (fs[8:15, 1] == fs[7, 1]).all()


def test_swing_trajectory_lands_on_target():
    """FootTrajectoryGenerator: apex = maxHeight at mid swing, z back to ~0 and (x, y) on the locked target at touch-down,
    velocity = finite difference of position (FootTrajectoryGenerator.cpp:41-108)."""
    P = ob.OraclePlanner(1)
    q, v, r = _idle_inputs(1)
    v[0, 0] = r[0, 0] = 0.3
    zs, xs, vx, tg = [], [], [], []
    for k in range(0, 160):
        P.run(k, q, v, r, stages=7)
        o = P.outputs()
        zs.append(o["fpos"][0, 5]), xs.append(o["fpos"][0, 3]), vx.append(o["fvel"][0, 3]), tg.append(o["o_target"][0, 3])
    zs, xs, vx = np.array(zs), np.array(xs), np.array(vx)
    # tick 0 rolls the gait, so FR swings during ticks 0..69 (t0 = 0.02 + 0.002 k, d = 0.16) and stands during 70..149
    assert abs(zs[:70].max() - 0.05) < 1e-3 and abs(zs[69]) < 1e-6 and (zs[70:150] == zs[69]).all()
    assert abs(xs[69] - tg[34]) < 1e-9  # target locked once t0 >= d - lockTime = 0.09, i.e. from tick 35 on
    fd = np.diff(xs[:70]) / 0.002
    assert np.abs(fd - 0.5 * (vx[:69] + vx[1:70])).max() < 5e-3
    assert zs[150] > 0 and (xs[70:150] == xs[69]).all()


# ------------------------------------------------------------------------------------------ emulation vs oracle (CPU)
@pytest.fixture(scope="module")
def emu():
    here = os.path.dirname(os.path.abspath(__file__))
    import subprocess

    subprocess.check_call(["make", "-C", os.path.join(here, "emu"), "-s"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    L = C.CDLL(os.path.join(here, "emu", "libqrw_emu.so"))
    L.emu_planner_create.restype = C.c_void_p
    L.emu_planner_create.argtypes = [C.c_int] + [C.c_double] * 4 + [C.c_int, C.c_int, C.c_double, _dp, _dp, _dp, C.c_double, C.c_double]
    L.emu_planner_destroy.argtypes = [C.c_void_p]
    L.emu_planner_run.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _dp, _ip]
    L.emu_planner_get.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
    L.emu_planner_gait.argtypes = [C.c_void_p, C.c_int, _dp]
    L.emu_planner_flags.argtypes = [C.c_void_p, _ip]
    L.emu_planner_field.argtypes = [C.c_char_p]
    return L


class EmuPlanner:
    def __init__(self, L, B, h_ref=H_REF):
        self.L, self.B = L, B
        sh = np.ascontiguousarray(ob.SHOULDERS34)
        self.h = L.emu_planner_create(B, 0.02, 0.002, 0.32, 0.32, 20, 10, h_ref, _p(sh), _p(sh), _p(sh), 0.05, 0.07)

    def __del__(self):
        self.L.emu_planner_destroy(self.h)

    def run(self, k, q7, hv, vr, code, stages=31):
        code = np.ascontiguousarray(code, np.int32)
        self.L.emu_planner_run(self.h, k, stages, _p(q7), _p(hv), _p(vr), code.ctypes.data_as(_ip))

    def get(self, name):
        B = self.B
        if name in ("xref", "fsteps"):
            out = np.empty((B, 204 if name == "xref" else 240))
            self.L.emu_planner_get(self.h, 0 if name == "xref" else 1, 0, _p(out))
        elif name in ("past", "gait", "desired"):
            out = np.empty((B, 80))
            self.L.emu_planner_gait(self.h, ("past", "gait", "desired").index(name), _p(out))
        else:
            n = 4 if name in ("t0s", "t_swing", "contacts") else 12
            out = np.empty((B, n))
            self.L.emu_planner_get(self.h, self.L.emu_planner_field(name.encode()), n, _p(out))
        return out

    def flags(self):
        f = np.empty((self.B, 2), np.int32)
        self.L.emu_planner_flags(self.h, f.ctypes.data_as(_ip))
        return f


def _compare_tick(get, flags, O, pv, q7, vr, tol):
    """One tick's outputs and state of an implementation (get(name), flags()) against the oracle O; pv = previous
    (p_cmd, v_cmd) for the feet commands, which the oracle side computes with numpy (Controller.py:286-296)."""
    o, s = O.outputs(), O.state()
    for n in ("gait",):
        assert (get(n) == o[n]).all(), n
    assert (get("past") == s["past"]).all() and (get("desired") == s["desired"]).all()
    assert (flags() == s["flags"]).all()
    assert (get("contacts") == o["gait"].reshape(-1, 4, 20)[:, :, 0]).all()
    for n in ("xref", "fsteps", "o_target", "fpos", "fvel", "facc"):
        assert _close(get(n), o[n], tol), (n, np.abs(get(n) - o[n]).max())
    assert ((get("fsteps") == 0.0) == (o["fsteps"] == 0.0)).all()  # swing entries exactly zero on both sides
    for n in ("cur_fs", "t0s", "t_swing"):
        assert _close(get(n), s[n], tol), n
    ref = ob.feet_commands(o["fpos"], o["fvel"], o["facc"], q7, vr, pv[0], pv[1], H_REF)
    for n, r in zip(("p_cmd", "v_cmd", "a_cmd"), ref):
        assert _close(get(n), r, tol), (n, np.abs(get(n) - r).max())
    return ref[0], ref[1]


@pytest.mark.parametrize("seed,codes", [(1, (1, 2, 3, 4)), (2, (3, 5)), (3, ())])
def test_emulated_planner_core_matches_oracle(emu, seed, codes):
    B, T = 24, 420
    W = wl.PlannerWorkload(B, seed, codes=codes, change_every=70)
    E, O = EmuPlanner(emu, B), ob.OraclePlanner(B, h_ref=H_REF)
    pv = (np.zeros((B, 12)), np.zeros((B, 12)))
    for k in range(T):
        q7, hv, vr, code = W.tick(k)
        E.run(k, q7, hv, vr, code)
        O.run(k, q7, hv, vr, code)
        pv = _compare_tick(E.get, E.flags, O, pv, q7, vr, 1e-11)


def test_emulated_stages_one_by_one_equal_the_fused_tick(emu):
    B = 8
    W = wl.PlannerWorkload(B, 5, change_every=50)
    A, S = EmuPlanner(emu, B), EmuPlanner(emu, B)
    for k in range(130):
        q7, hv, vr, code = W.tick(k)
        A.run(k, q7, hv, vr, code)
        for st in (1, 2, 4, 8, 16):
            S.run(k, q7, hv, vr, code, stages=st)
        for n in ("xref", "fsteps", "gait", "past", "desired") + FIELDS:
            assert (A.get(n) == S.get(n)).all(), (k, n)


def test_emulated_planner_does_not_depend_on_the_lane_order(emu):
    """The four lanes of a robot exchange data only through the __syncwarp between lane_compute and lane_store: any order
    of the lanes inside a phase must give the same bits (race check without a GPU)."""
    B = 12
    W = wl.PlannerWorkload(B, 7, codes=(1, 2, 3, 4), change_every=40)
    P = [EmuPlanner(emu, B) for _ in range(3)]
    try:
        for k in range(170):
            q7, hv, vr, code = W.tick(k)
            for order, p in enumerate(P):
                emu.emu_planner_set_lane_order(order)
                p.run(k, q7, hv, vr, code)
            for n in ("xref", "fsteps", "gait", "past", "desired") + FIELDS:
                assert (P[0].get(n) == P[1].get(n)).all() and (P[0].get(n) == P[2].get(n)).all(), (k, n)
    finally:
        emu.emu_planner_set_lane_order(0)


# ------------------------------------------------------------------------------------------ CUDA vs oracle (GPU)
@pytest.mark.gpu
@pytest.mark.parametrize("B,seed,codes", [(37, 1, (1, 2, 3, 4)), (256, 2, (3, 5))])
def test_cuda_planner_matches_oracle(B, seed, codes):
    from qrw_b200.planner import BatchedPlanner

    W = wl.PlannerWorkload(B, seed, codes=codes, change_every=70)
    G, O = BatchedPlanner(B, h_ref=H_REF), ob.OraclePlanner(B, h_ref=H_REF)
    pv = (np.zeros((B, 12)), np.zeros((B, 12)))
    for k in range(300):
        q7, hv, vr, code = W.tick(k)
        G.run(k, q7, hv, vr, code)
        O.run(k, q7, hv, vr, code)
        out, st = G.outputs(), G.state()
        tens = None

        def get(n):
            if n in out:
                return out[n]
            if n == "contacts":
                return out["gait"].reshape(-1, 4, 20)[:, :, 0]
            return st[n]

        pv = _compare_tick(get, lambda: st["flags"], O, pv, q7, vr, 1e-9)


@pytest.mark.gpu
def test_cuda_planner_device_path_and_full_size_properties():
    """65,536 robots through the device-pointer entry point: equals the host entry point bit for bit, the gait read back
    from fsteps (MPC::construct_gait's rule) equals the gait matrix, contacts = row 0."""
    import torch

    from qrw_b200.planner import BatchedPlanner

    B = 65536
    W = wl.PlannerWorkload(B, 4, codes=(1, 2, 3, 4, 5), change_every=40)
    H, D = BatchedPlanner(B), BatchedPlanner(B)
    for k in range(101):
        q7, hv, vr, code = W.tick(k)
        H.run(k, q7, hv, vr, code)
        D.run(k, torch.from_numpy(q7).cuda(), torch.from_numpy(hv).cuda(), torch.from_numpy(vr).cuda(), torch.from_numpy(code).cuda())
    torch.cuda.synchronize()
    oh, od, t = H.outputs(), D.outputs(), D.output_tensors()
    for n in oh:
        assert (oh[n] == od[n]).all(), n
    g = oh["gait"].reshape(B, 4, 20)
    fs = oh["fsteps"].reshape(B, 4, 3, 20)
    assert ((fs[:, :, 0, :] != 0.0) == (g == 1.0)).all()
    assert (t["contacts"].cpu().numpy() == g[:, :, 0]).all()
    assert (t["xref"].cpu().numpy() == oh["xref"]).all() and (t["fsteps"].cpu().numpy() == oh["fsteps"]).all()
    assert np.isfinite(oh["fpos"]).all() and np.isfinite(oh["a_cmd"]).all()


@pytest.mark.gpu
def test_cuda_planner_rejects_small_gait_matrix():
    from qrw_b200 import QrwError
    from qrw_b200.planner import BatchedPlanner

    with pytest.raises(QrwError, match="Increase N_gait"):
        BatchedPlanner(4, N_gait=12)


@pytest.mark.gpu
def test_drop_in_planner_classes_follow_the_controller_call_sequence():
    """The Boost.Python-shaped classes (libquadruped_reactive_walking.Gait / StatePlanner / FootstepPlanner /
    FootTrajectoryGenerator) driven exactly as scripts/Controller.py:124-139,222-240 drives them, against the oracle."""
    import libquadruped_reactive_walking as lqrw

    dt_wbc, dt_mpc, T, N_gait, k_mpc = 0.002, 0.02, 0.32, 20, 10
    sh = ob.SHOULDERS34.reshape(4, 3).T.copy()
    sp, gait, fsp, ftg = lqrw.StatePlanner(), lqrw.Gait(), lqrw.FootstepPlanner(), lqrw.FootTrajectoryGenerator()
    sp.initialize(dt_mpc, T, H_REF)
    gait.initialize(dt_mpc, T, T, N_gait)
    fsp.initialize(dt_mpc, dt_wbc, T, H_REF, sh.copy(), gait, N_gait)
    ftg.initialize(0.05, 0.07, sh.copy(), sh.copy(), dt_wbc, k_mpc, gait)
    O = ob.OraclePlanner(1, h_ref=H_REF)
    W = wl.PlannerWorkload(1, 31, codes=(1, 2, 3), change_every=45)
    for k in range(140):
        q7, hv, vr, code = W.tick(k)
        q = np.zeros((19, 1))
        q[:7, 0] = q7[0]
        gait.updateGait(k, k_mpc, q[0:7, 0:1], int(code[0]))
        o_t = fsp.updateFootsteps(k % k_mpc == 0 and k != 0, int(k_mpc - k % k_mpc), q[0:7, 0:1], hv[0].reshape(6, 1).copy(), vr[0])
        ftg.update(k, o_t)
        sp.computeReferenceStates(q[0:7, 0:1], hv[0].reshape(6, 1).copy(), vr[0].reshape(6, 1), 0.0)
        O.run(k, q7, hv, vr, code)
        o = O.outputs()
        assert (gait.getCurrentGait() == o["gait"].reshape(4, 20).T).all()
        assert gait.isNewPhase() == bool(O.state()["flags"][0, 0])
        assert _close(sp.getReferenceStates(), o["xref"].reshape(17, 12).T)
        assert _close(fsp.getFootsteps(), o["fsteps"].reshape(12, 20).T)
        assert _close(o_t, o["o_target"].reshape(4, 3).T)
        for got, name in ((ftg.getFootPosition(), "fpos"), (ftg.getFootVelocity(), "fvel"), (ftg.getFootAcceleration(), "facc")):
            assert _close(got, o[name].reshape(4, 3).T), (k, name)
    with pytest.raises(ValueError):
        lqrw.Gait().initialize(dt_mpc, T, T, 12)
