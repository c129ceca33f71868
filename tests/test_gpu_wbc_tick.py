"""GPU whole-body tick (Solo12 kinematics / RNEA + InvKin + QPWBC + tau_ff, csrc/solo12_core.cuh) against the oracle's
composition of the same steps (oracle.bindings.wbc_tick), BASELINE config 4 inputs."""
import numpy as np
import pytest

from conftest import assert_forces_close

pytestmark = pytest.mark.gpu


def test_foot_kinematics_match_oracle(oracle):
    from qrw_b200 import solo12_feet

    rng = np.random.default_rng(0)
    q, dq = rng.uniform(-1.2, 1.2, (257, 12)), rng.uniform(-4, 4, (257, 12))
    for got, want in zip(solo12_feet(q, dq), oracle.solo12_feet(q, dq)):
        np.testing.assert_allclose(got, want, rtol=1e-11, atol=1e-11)


def test_base_inertia_matches_crba(oracle):
    from qrw_b200 import solo12_base_inertia

    q0 = np.zeros(19)
    q0[6] = 1.0
    np.testing.assert_allclose(solo12_base_inertia(), np.diag(oracle.solo12_crba(q0))[:6], rtol=1e-12)


@pytest.mark.parametrize("level,group", [(True, 0), (False, 8), (False, 4), (True, 2)])
def test_wbc_tick_matches_oracle(oracle, level, group):
    """B = 301: the last group of 8 / 4 / 2 robots is partly empty; group 0 = the library's own choice."""
    from qrw_b200 import BatchedWbcTick, workloads as wl

    B = 301
    gpu, cpu = BatchedWbcTick(B, debug=True, group=group), oracle.OracleWBC(B)
    for call in range(3):  # the QP warm start carries over
        args = wl.wbc_tick_inputs(B, seed=40 + call, level_base=level)
        gpu.compute(*args)
        ref = oracle.wbc_tick(cpu, *args, nthreads=8)
        dbg = gpu.debug()
        np.testing.assert_allclose(dbg["Jc"], ref["Jc"], rtol=1e-11, atol=1e-11)
        np.testing.assert_allclose(dbg["RNEA"], ref["RNEA"], rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(gpu.ddq_cmd, ref["ddq_cmd"], rtol=1e-9, atol=1e-8)
        np.testing.assert_allclose(gpu.qdes, ref["qdes"], rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(gpu.vdes, ref["vdes"], rtol=1e-10, atol=1e-10)
        assert_forces_close(gpu.f_with_delta, ref["f_res"], f"f_with_delta call {call}")
        np.testing.assert_allclose(gpu.tau_ff, ref["tau_ff"], rtol=1e-5, atol=1e-4)


def test_wbc_tick_does_not_depend_on_the_group_size():
    """Who shares a robot's warp changes nothing: groups of 8, 4 and 2 give the same bits (two calls, warm start)."""
    from qrw_b200 import BatchedWbcTick, workloads as wl

    B = 1021
    res = []
    for group in (8, 4, 2):
        gpu = BatchedWbcTick(B, group=group)
        got = []
        for call in range(2):
            gpu.compute(*wl.wbc_tick_inputs(B, seed=60 + call, level_base=False))
            got.append([a.copy() for a in (gpu.tau_ff, gpu.qdes, gpu.vdes, gpu.f_with_delta, gpu.ddq_cmd)])
        res.append(got)
    for other in res[1:]:
        for a, b in zip(res[0], other):
            for x, y in zip(a, b):
                assert np.array_equal(x, y)


def test_wbc_tick_full_size_properties():
    """BASELINE config 4 size (65536 robots): swing feet get no force, stance forces stay in the friction cone."""
    from qrw_b200 import BatchedWbcTick, workloads as wl

    B = 65536
    args = wl.wbc_tick_inputs(B, seed=3)
    gpu = BatchedWbcTick(B)
    gpu.compute(*args)
    f = gpu.f_with_delta.reshape(B, 4, 3)
    contacts = args[3]
    assert np.abs(f[contacts == 0]).max() < 1e-6
    fz = f[..., 2]
    assert (fz > -1e-3).all() and (fz < 25 + 1e-3).all()
    assert (np.abs(f[..., 0]) <= 0.9 * fz + 1e-3).all() and (np.abs(f[..., 1]) <= 0.9 * fz + 1e-3).all()
    assert np.isfinite(gpu.tau_ff).all()


def test_wbc_tick_full_size_sampled_against_oracle(oracle):
    """BASELINE config 4 at its full size (65,536 robots, groups of eight per warp, every CTA slot of the GPU busy): 256 robots
    sampled across the batch are recomputed by the oracle from the same inputs (cold start: robots are independent)."""
    from qrw_b200 import BatchedWbcTick, workloads as wl

    B = 65536
    args = wl.wbc_tick_inputs(B, seed=9, level_base=False)
    gpu = BatchedWbcTick(B)
    gpu.compute(*args)
    pick = np.unique(np.concatenate([np.arange(0, B, 257), [B - 1, B - 2, B - 9]]))
    sub = [np.ascontiguousarray(a[pick]) for a in args]
    ref = oracle.wbc_tick(oracle.OracleWBC(len(pick)), *sub, nthreads=8)
    assert_forces_close(gpu.f_with_delta[pick], ref["f_res"], "f_with_delta, sampled robots")
    np.testing.assert_allclose(gpu.tau_ff[pick], ref["tau_ff"], rtol=1e-5, atol=1e-4)
    np.testing.assert_allclose(gpu.qdes[pick], ref["qdes"], rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(gpu.vdes[pick], ref["vdes"], rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(gpu.ddq_cmd[pick], ref["ddq_cmd"], rtol=1e-9, atol=1e-8)
