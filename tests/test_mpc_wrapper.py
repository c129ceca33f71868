"""Host logic of ``BatchedMPCWrapper`` (reference scripts/MPC_Wrapper.py:35-148 for a batch): default first result, pass-through
of ``solve`` -- with a stub solver on CPU and the CUDA solver on the GPU.  (The single-robot wrapper is the reference's own script
on the drop-in module: tests/test_reference_scripts.py.)"""
import numpy as np
import pytest


class _StubBatchedMPC:
    def __init__(self, B):
        self.B, self.calls = B, []

    def run(self, k, xref, fsteps):
        self.calls.append((k, np.array(xref, copy=True), np.array(fsteps, copy=True)))
        return 0

    def get_latest_result(self):
        return np.full((self.B, 384), float(len(self.calls)))


def _q_init(B):
    q = np.zeros((B, 19))
    q[:, 2], q[:, 6] = 0.2447495, 1.0
    return q


def test_first_result_then_solver_results():
    from qrw_b200.mpc_wrapper import BatchedMPCWrapper

    B = 3
    w = BatchedMPCWrapper(B, 0.02, 16, 10, 0.32, 20, _q_init(B), mpc=_StubBatchedMPC(B))
    first = w.get_latest_result()
    assert first.shape == (B, 24, 16) and np.allclose(first[:, 12:, 0], [0, 0, 8] * 4) and (first[:, 2, 0] == 0.2447495).all()
    assert not first[:, :, 1:].any()
    w.solve(0, np.zeros((B, 204)), np.zeros((B, 240)))
    got = w.get_latest_result()
    assert got.shape == (B, 24, 16) and (got == 1.0).all()
    w.solve(10, np.zeros((B, 204)), np.zeros((B, 240)))
    assert (w.get_latest_result() == 2.0).all()
    assert w.mpc.calls[1][0] == 10 and isinstance(w.mpc.calls[1][0], int)


@pytest.mark.gpu
def test_wrapper_on_the_cuda_solver_matches_oracle(oracle):
    """scripts/test_mpc.py:136-160 shape: trot, three MPC calls through the wrapper; forces equal the oracle's."""
    from conftest import assert_forces_close
    from qrw_b200 import workloads as wl
    from qrw_b200.mpc_wrapper import BatchedMPCWrapper

    w = BatchedMPCWrapper(1, 0.02, 16, 10, 0.32, 20, _q_init(1))
    o = oracle.OracleMPC(1)
    xr, fs = wl.config1_single_trot()
    assert np.allclose(w.get_latest_result()[0, 12:, 0], [0, 0, 8] * 4)
    for k in (0, 10, 20):
        w.solve(k, xr, fs)
        o.run(k, xr, fs)
        got = w.get_latest_result()
        assert got.shape == (1, 24, 16)
        assert_forces_close(got[0], o.get_latest_result().reshape(16, 24).T, f"k={k}")
