"""Host logic of the MPC_Wrapper mirror (reference scripts/MPC_Wrapper.py:35-148): default first result, roll and
refill of last_available_result, first-call behaviour -- with a stub solver on CPU and the CUDA solver on the GPU."""
import numpy as np
import pytest


class _StubMPC:
    def __init__(self):
        self.calls = []

    def run(self, k, xref, fsteps):
        self.calls.append((k, xref.copy(), fsteps.copy()))
        return 0

    def get_latest_result(self):
        return np.full((24, 16), float(len(self.calls)))


def _q_init():
    q = np.zeros((19, 1))
    q[2, 0], q[6, 0] = 0.2447495, 1.0
    return q


def test_first_result_roll_and_refill():
    from qrw_b200.mpc_wrapper import MPC_Wrapper

    w = MPC_Wrapper(True, 0.02, 16, 10, 0.32, 20, _q_init(), False, mpc=_StubMPC())
    first = w.get_latest_result()
    assert first.shape == (24, 16) and np.allclose(first[12:, 0], [0, 0, 8] * 4) and first[2, 0] == 0.2447495
    assert not first[:, 1:].any()
    gait = np.zeros((20, 4))
    gait[:8] = [1, 0, 0, 1]
    gait[8:16] = [0, 1, 1, 0]
    xref, fsteps = np.zeros((12, 17)), np.zeros((20, 12))
    w.solve(0, xref, fsteps, gait)                      # k <= 2: no roll, no refill
    assert np.allclose(w.last_available_result[12:, 0], [0, 0, 8] * 4)
    assert np.array_equal(w.get_latest_result(), np.full((24, 16), 1.0))
    w.solve(10, xref, fsteps, gait)                     # roll left, refill the last column from the last active row
    assert not w.last_available_result[12:, 0].any()    # the default column was rolled out
    last = w.last_available_result[12:, 15].reshape(4, 3)
    assert np.allclose(last[[1, 2], 2], 9.81 * 2.5 / 2) and not last[[0, 3]].any()
    assert w.mpc.calls[1][0] == 10 and isinstance(w.mpc.calls[1][0], int)


def test_unsupported_modes_are_loud():
    from qrw_b200.mpc_wrapper import MPC_Wrapper

    with pytest.raises(NotImplementedError):
        MPC_Wrapper(True, 0.02, 16, 10, 0.32, 20, _q_init(), True, mpc=_StubMPC())
    with pytest.raises(NotImplementedError):
        MPC_Wrapper(False, 0.02, 16, 10, 0.32, 20, _q_init(), False, mpc=_StubMPC())


@pytest.mark.gpu
def test_wrapper_on_the_cuda_solver_matches_oracle(oracle):
    """scripts/test_mpc.py:136-160 shape: trot, closed loop through the wrapper; forces equal the oracle's."""
    from conftest import assert_forces_close
    from qrw_b200 import workloads as wl
    from qrw_b200.mpc_wrapper import MPC_Wrapper

    w = MPC_Wrapper(True, 0.02, 16, 10, 0.32, 20, _q_init(), False)
    o = oracle.OracleMPC(1)
    xr, fs = wl.config1_single_trot()
    xref, fsteps = xr.reshape(17, 12).T.copy(), fs.reshape(12, 20).T.copy()
    gait = (fsteps[:, ::3] != 0).astype(float)
    assert np.allclose(w.get_latest_result()[12:, 0], [0, 0, 8] * 4)
    for k in (0, 10, 20):
        w.solve(k, xref, fsteps, gait)
        o.run(k, xr, fs)
        got = w.get_latest_result()
        assert got.shape == (24, 16)
        assert_forces_close(got, o.get_latest_result().reshape(16, 24).T, f"k={k}")
