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
