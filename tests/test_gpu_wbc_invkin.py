"""GPU parity of QPWBC (src/QPWBC.cpp:310-343 + OSQP) and InvKin (src/InvKin.cpp:23-73)."""
import numpy as np
import pytest

from conftest import assert_forces_close

pytestmark = pytest.mark.gpu


def test_wbc_matches_oracle_over_three_calls(oracle):
    from qrw_b200 import BatchedQPWBC, workloads as wl

    B = 256
    gpu, cpu = BatchedQPWBC(B), oracle.OracleWBC(B)
    for call in range(3):  # first call = osqp_setup, later ones = update_P/_lin_cost/_bounds, warm start
        M, Jc, f, R = wl.wbc_inputs(B, seed=30 + call)
        gpu.run(M, Jc, f, R)
        cpu.run(M, Jc, f, R, 8)
        fg, dg, Hg = gpu.results()
        fc, dc, Hc = cpu.results()
        assert np.array_equal(gpu.info()["iters"], cpu.info()["iters"])
        assert_forces_close(fg, fc, f"f_res call {call}")
        np.testing.assert_allclose(dg, dc, rtol=1e-5, atol=1e-4)
        np.testing.assert_allclose(Hg, Hc, rtol=1e-12, atol=1e-12)


def test_invkin_matches_oracle(oracle):
    from qrw_b200 import BatchedInvKin, workloads as wl

    B = 1000
    args = wl.invkin_inputs(B, seed=4)
    ik = BatchedInvKin(B)
    ik.refreshAndCompute(*args)
    ddq, dq, qs = oracle.invkin(*args)
    np.testing.assert_allclose(ik.ddq, ddq, rtol=1e-11, atol=1e-9)
    np.testing.assert_allclose(ik.get_dq_cmd(), dq, rtol=1e-11, atol=1e-11)
    np.testing.assert_allclose(ik.get_q_step(), qs, rtol=1e-11, atol=1e-12)


def test_drop_in_module_classes(oracle):
    """The Boost.Python-shaped module (python/gepadd.cpp:178-238): one robot, natural (rows x cols) numpy matrices in,
    the reference's getter shapes out."""
    import libquadruped_reactive_walking as lrw
    from qrw_b200 import workloads as wl
    from qrw_b200.batched import unpack_cols

    M, Jc, f, R = wl.wbc_inputs(1, seed=77)
    q = lrw.QPWBC()
    assert q.run(unpack_cols(M, 18, 18)[0], unpack_cols(Jc, 12, 18)[0], f.reshape(12, 1), R.reshape(6, 1), np.zeros((1, 4))) == 0
    o = oracle.OracleWBC(1)
    o.run(M, Jc, f, R)
    fc, dc, Hc = o.results()
    assert q.get_f_res().shape == (12, 1) and q.get_ddq_res().shape == (6, 1) and q.get_H().shape == (12, 12)
    assert_forces_close(q.get_f_res().ravel(), fc.ravel(), "QPWBC.get_f_res")
    np.testing.assert_allclose(q.get_H(), Hc.reshape(12, 12).T, rtol=1e-12, atol=1e-12)
    args = wl.invkin_inputs(1, seed=78)
    shapes = [(1, 4), (3, 4), (3, 4), (3, 4), (4, 3), (4, 3), (4, 3), (4, 3), (12, 12)]
    nat = [unpack_cols(a, r, c)[0] for a, (r, c) in zip(args, shapes)]
    ik = lrw.InvKin(0.002)
    ddq = ik.refreshAndCompute(*nat)
    ref = oracle.invkin(*args)
    assert ddq.shape == (12, 1) and ik.get_q_step().shape == (12, 1)
    np.testing.assert_allclose(ddq.ravel(), ref[0].ravel(), rtol=1e-11, atol=1e-9)
    np.testing.assert_allclose(ik.get_dq_cmd().ravel(), ref[1].ravel(), rtol=1e-11, atol=1e-11)
