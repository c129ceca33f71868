"""The product's kernel cores (csrc/mpc_core.cuh, wbc_core.cuh), compiled by g++ in warp-emulation
mode (tests/emu), against the CPU oracle on the same seeded inputs and the same call history.
This is the no-GPU check of the CUDA logic; the `-m gpu` tests repeat it on the real kernels."""
import ctypes as C

import numpy as np
import pytest

from conftest import assert_forces_close

GAITS = ("trot", "walk", "bounding", "pacing", "static")


from emu_backend import EmuMPC, _p  # noqa: E402


@pytest.mark.parametrize("core", ["v2", "v4", "v3", "cta", "warp"])
@pytest.mark.parametrize("gaits,seed", [(("trot",), 1), (GAITS, 2)])
def test_mpc_closed_loop(emu, oracle, gaits, seed, core):
    from qrw_b200 import workloads as wl

    B, ticks = 24, 4
    w = wl.MpcWorkload(B, seed, gaits=gaits, ticks=ticks + 1)
    e, o = EmuMPC(emu, B, core=core), oracle.OracleMPC(B)
    x0 = None
    for t in range(ticks):
        xr, fs = w.inputs(t, x0)
        e.run(t, xr, fs)
        o.run(t, xr, fs, 8)
        ro, io = o.get_latest_result(), o.info()
        assert np.array_equal(e.gait(), o.get_gait())
        assert np.array_equal(e.sgait(), o.get_Sgait())
        assert np.array_equal(e.info[:, 0], io["iters"]), f"tick {t}"
        assert np.array_equal(e.info[:, 1], io["status"])
        assert np.array_equal(e.info[:, 2], io["n_factor"] - (0 if t == 0 else _prev_nf)), "KKT factorisations"
        _prev_nf = io["n_factor"].copy()
        np.testing.assert_allclose(e.rho, io["rho"], rtol=1e-5)  # sqrt of a ratio of 1e-6-sized residuals: round-off is amplified
        assert_forces_close(e.out, ro, f"x_f_applied tick {t}")
        assert np.abs(e.out - ro).max() < 1e-7  # in practice round-off only
        x0 = ro[:, :12].copy()


@pytest.mark.parametrize("core", ["v2", "v4", "v3", "cta"])
def test_mpc_result_does_not_depend_on_the_thread_schedule(emu, core):
    """Race check without a GPU: between two collectives the emulator may run the threads of a CTA in any order (every order
    is a legal CUDA schedule).  The chain / helper warps of the v2 core hand data over through named barriers; a missing
    barrier shows up as a result that depends on the order."""
    from qrw_b200 import workloads as wl

    B, ticks = 6, 3
    w = wl.MpcWorkload(B, 5, gaits=GAITS, ticks=ticks + 1)
    runs = []
    try:
        for order in (0, 1, 2, 3, 4):
            emu.emu_set_schedule(order)
            e, x0, outs = EmuMPC(emu, B, core=core), None, []
            for t in range(ticks):
                xr, fs = w.inputs(t, x0)
                e.run(t, xr, fs)
                outs.append((e.out.copy(), e.info.copy(), e.std.copy()))
                x0 = e.out.reshape(B, 16, 24)[:, 0, :12].copy()
            runs.append(outs)
    finally:
        emu.emu_set_schedule(0)
    for other in runs[1:]:
        for (o0, i0, s0), (o1, i1, s1) in zip(runs[0], other):
            assert np.array_equal(i0, i1)
            assert np.array_equal(o0, o1) and np.array_equal(s0, s1)  # bit for bit


@pytest.mark.parametrize("core", ["v2", "cta", "warp"])
def test_mpc_scaling_and_bounds(emu, oracle, core):
    from qrw_b200 import workloads as wl

    B = 6
    w = wl.MpcWorkload(B, 3, gaits=GAITS, ticks=3)
    e, o = EmuMPC(emu, B, core=core), oracle.OracleMPC(B)
    x0 = None
    for t in range(2):
        xr, fs = w.inputs(t, x0)
        e.run(t, xr, fs)
        o.run(t, xr, fs, 4)
        for b in range(B):
            s = o.solver_state(b)
            np.testing.assert_allclose(e.dbg[b, :384], s["D"], rtol=1e-12)
            np.testing.assert_allclose(e.dbg[b, 384:1088], s["E"], rtol=1e-12)
            np.testing.assert_allclose(e.dbg[b, 1088], s["c"], rtol=1e-12)
            np.testing.assert_allclose(e.dbg[b, 1089:], o.bounds(b)[1][:192], rtol=0, atol=1e-15)
        x0 = o.get_latest_result()[:, :12].copy()


@pytest.mark.parametrize("core", ["v2", "cta", "warp"])
def test_mpc_flight_phase_and_stale_rows(emu, oracle, core):
    """Quirk A.6-2: an all-swing fsteps row terminates gait processing; later steps keep the stale
    B / S values of earlier calls."""
    from qrw_b200 import workloads as wl

    xr, fs = wl.config1_single_trot()
    fs2 = fs.reshape(12, 20).copy()
    fs2[:, 10:] = 0.0  # only 10 active rows on the second call
    e, o = EmuMPC(emu, 1, core=core), oracle.OracleMPC(1)
    for k, f in ((0, fs), (1, fs2.reshape(1, -1)), (2, fs)):
        e.run(k, xr, f, 1)
        o.run(k, xr, f)
        assert np.array_equal(e.gait(), o.get_gait())
        assert np.array_equal(e.sgait(), o.get_Sgait())
        assert e.info[0, 0] == o.info()["iters"][0]
        assert_forces_close(e.out, o.get_latest_result(), f"run({k})")


@pytest.mark.parametrize("core", ["v2", "cta", "warp"])
def test_mpc_max_iter_returns_last_iterate(emu, oracle, core):
    """Quirk A.6-8: on max_iter the last iterate is returned (status -2 / 2), never an error."""
    from qrw_b200 import workloads as wl

    xr, fs = wl.config1_single_trot()
    e, o = EmuMPC(emu, 1, core=core), oracle.OracleMPC(1)
    o.set_max_iter(50)
    e.run(0, xr, fs, 1, max_iter=50)
    o.run(0, xr, fs)
    assert e.info[0, 0] == 50 and o.info()["iters"][0] == 50
    assert e.info[0, 1] == o.info()["status"][0]
    np.testing.assert_allclose(e.out, o.get_latest_result(), rtol=1e-9, atol=1e-9)


def test_wbc_three_calls(emu, oracle):
    from qrw_b200 import workloads as wl

    B = 64
    o = oracle.OracleWBC(B)
    st = np.zeros((B, emu.emu_wbc_state_doubles()))
    f_res, ddq, H, info = np.zeros((B, 12)), np.zeros((B, 6)), np.zeros((B, 144)), np.zeros((B, 4), np.int32)
    for call in range(3):
        M, Jc, f, R = wl.wbc_inputs(B, seed=30 + call)
        emu.emu_wbc_run(B, int(call == 0), _p(M), _p(Jc), _p(f), _p(R), _p(st), _p(f_res), _p(ddq), _p(H), _p(info))
        o.run(M, Jc, f, R, 4)
        fc, dc, Hc = o.results()
        assert np.array_equal(info[:, 0], o.info()["iters"]), f"call {call}"
        assert_forces_close(f_res, fc, f"f_res call {call}")
        np.testing.assert_allclose(ddq, dc, rtol=1e-5, atol=1e-4)
        np.testing.assert_allclose(H, Hc, rtol=1e-12, atol=1e-12)


def test_warp_cores_do_not_depend_on_the_lane_order(emu):
    """wbc_core.cuh / mpc_core.cuh run the 32 lanes of a phase one after the other in emulation; a phase that reads what
    another lane writes in the same phase (a missing __syncwarp on the GPU) gives different bits with the lanes reversed."""
    from qrw_b200 import workloads as wl

    B = 16
    res = []
    try:
        for desc in (0, 1):
            emu.emu_set_lanes_descending(desc)
            st = np.zeros((B, emu.emu_wbc_state_doubles()))
            f_res, ddq, H, info = np.zeros((B, 12)), np.zeros((B, 6)), np.zeros((B, 144)), np.zeros((B, 4), np.int32)
            got = []
            for call in range(3):
                M, Jc, f, R = wl.wbc_inputs(B, seed=30 + call)
                emu.emu_wbc_run(B, int(call == 0), _p(M), _p(Jc), _p(f), _p(R), _p(st), _p(f_res), _p(ddq), _p(H), _p(info))
                got.append((f_res.copy(), ddq.copy(), H.copy(), info.copy(), st.copy()))
            w = wl.MpcWorkload(4, 3, gaits=GAITS, ticks=3)
            e, x0 = EmuMPC(emu, 4, core="warp"), None
            for t in range(2):
                xr, fs = w.inputs(t, x0)
                e.run(t, xr, fs)
                got.append((e.out.copy(), e.info.copy(), e.std.copy()))
                x0 = e.out.reshape(4, 16, 24)[:, 0, :12].copy()
            res.append(got)
    finally:
        emu.emu_set_lanes_descending(0)
    for a, b in zip(res[0], res[1]):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)


def test_invkin(emu, oracle):
    from qrw_b200 import workloads as wl

    B = 50
    args = wl.invkin_inputs(B, seed=4)
    outs = [np.zeros((B, 12)) for _ in range(3)]
    emu.emu_invkin_run(B, *[_p(a) for a in args], *[_p(o) for o in outs])
    for a, b in zip(outs, oracle.invkin(*args)):
        np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-12)


def _emu_wbc_tick(emu, B, G, first, args, st, M):
    q, dq, f, contacts, pg, vg, ag = [np.ascontiguousarray(a, np.float64) for a in args]
    out = dict(f_res=np.zeros((B, 12)), ddq_res=np.zeros((B, 6)), info=np.zeros((B, 4), np.int32), ddq_cmd=np.zeros((B, 12)),
               vdes=np.zeros((B, 12)), qdes=np.zeros((B, 12)), tau_ff=np.zeros((B, 12)), Jc=np.zeros((B, 216)), RNEA=np.zeros((B, 6)))
    emu.emu_wbc_tick_run(B, G, int(first), _p(q), _p(dq), _p(f), 12, _p(contacts), _p(pg), _p(vg), _p(ag), _p(M), _p(st),
                         _p(out["f_res"]), _p(out["ddq_res"]), _p(out["info"]), _p(out["ddq_cmd"]), _p(out["vdes"]), _p(out["qdes"]),
                         _p(out["tau_ff"]), _p(out["Jc"]), _p(out["RNEA"]))
    return out


def _wbc_tick_M(oracle):
    q0 = np.zeros(19)
    q0[6] = 1.0
    M = oracle.solo12_crba(q0)
    M[:6, :6] = M[:6, :6] * np.eye(6)  # scripts/QP_WBC.py:93
    return np.asfortranarray(M).ravel(order="F").copy()


@pytest.mark.parametrize("G,B", [(8, 37), (4, 22), (2, 9)])
def test_wbc_tick_groups_three_calls(emu, oracle, G, B):
    """csrc/wbc_tick_core.cuh (eight / four / two robots per warp, the QP on half-warps) in the fiber emulation against the
    oracle's composition of wbc_controller.compute; ragged batches (the last group is partly empty), warm start over three calls."""
    from qrw_b200 import workloads as wl

    cpu = oracle.OracleWBC(B)
    st = np.zeros((B, emu.emu_wbc_state_doubles()))
    M = _wbc_tick_M(oracle)
    for call in range(3):
        args = wl.wbc_tick_inputs(B, seed=40 + call, level_base=(call != 1))
        got = _emu_wbc_tick(emu, B, G, call == 0, args, st, M)
        ref = oracle.wbc_tick(cpu, *args, nthreads=4)
        np.testing.assert_allclose(got["Jc"], ref["Jc"], rtol=1e-11, atol=1e-11)
        np.testing.assert_allclose(got["RNEA"], ref["RNEA"], rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(got["ddq_cmd"], ref["ddq_cmd"], rtol=1e-9, atol=1e-8)
        np.testing.assert_allclose(got["qdes"], ref["qdes"], rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(got["vdes"], ref["vdes"], rtol=1e-10, atol=1e-10)
        assert np.array_equal(got["info"][:, 0], cpu.info()["iters"]), f"call {call}"
        assert_forces_close(got["f_res"], ref["f_res"], f"f_res call {call}")
        np.testing.assert_allclose(got["ddq_res"], ref["ddq_res"], rtol=1e-5, atol=1e-4)
        np.testing.assert_allclose(got["tau_ff"], ref["tau_ff"], rtol=1e-5, atol=1e-4)


def test_wbc_tick_groups_do_not_depend_on_the_thread_schedule_or_the_group_size(emu, oracle):
    """Five legal schedules of the emulated threads and the three group sizes give the same bits: no race between the
    collectives, and a robot's result does not depend on who shares its warp."""
    from qrw_b200 import workloads as wl

    B = 19
    M = _wbc_tick_M(oracle)
    res = []
    try:
        for order, G in ((0, 8), (1, 8), (2, 8), (3, 4), (4, 2)):
            emu.emu_set_schedule(order)
            st = np.zeros((B, emu.emu_wbc_state_doubles()))
            got = []
            for call in range(2):
                o = _emu_wbc_tick(emu, B, G, call == 0, wl.wbc_tick_inputs(B, seed=50 + call, level_base=False), st, M)
                got.append([o[k].copy() for k in sorted(o)] + [st.copy()])
            res.append(got)
    finally:
        emu.emu_set_schedule(0)
    for other in res[1:]:
        for a, b in zip(res[0], other):
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
