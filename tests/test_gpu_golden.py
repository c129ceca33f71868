"""GPU path through the C ABI against the committed golden fixtures (tests/golden, tools/make_golden.py):
needs neither the oracle's sources nor /root/reference on the GPU box."""
import os

import numpy as np
import pytest

from conftest import ROOT, assert_forces_close

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.mark.parametrize("name,B", [("mpc_config1_single_trot", 1), ("mpc_trot_b8", 8), ("mpc_mixed_b10", 10)])
def test_mpc_golden(name, B):
    from qrw_b200 import BatchedMPC

    g = np.load(os.path.join(GOLD, name + ".npz"))
    m = BatchedMPC(B)
    for t in range(int(g["ticks"])):
        m.run(t, g[f"xref{t}"], g[f"fsteps{t}"])
        assert np.array_equal(m.get_gait(), g[f"gait{t}"]) and np.array_equal(m.get_Sgait(), g[f"sgait{t}"])
        assert np.array_equal(m.info()["iters"], g[f"iters{t}"])
        assert np.array_equal(m.info()["status"], g[f"status{t}"])
        assert_forces_close(m.get_latest_result(), g[f"result{t}"], f"{name} tick {t}")


def test_wbc_invkin_golden():
    from qrw_b200 import BatchedInvKin, BatchedQPWBC

    g = np.load(os.path.join(GOLD, "wbc_b16.npz"))
    q = BatchedQPWBC(16)
    for c in range(3):
        q.run(g[f"M{c}"], g[f"Jc{c}"], g[f"f_cmd{c}"], g[f"RNEA{c}"])
        f, d, H = q.results()
        assert np.array_equal(q.info()["iters"], g[f"iters{c}"])
        assert_forces_close(f, g[f"f_res{c}"], f"f_res call {c}")
        np.testing.assert_allclose(d, g[f"ddq_res{c}"], rtol=1e-5, atol=1e-4)
        np.testing.assert_allclose(H, g[f"H{c}"], rtol=1e-12, atol=1e-12)
    k = np.load(os.path.join(GOLD, "invkin_b16.npz"))
    names = ["contacts", "goals", "vgoals", "agoals", "posf", "vf", "wf", "af", "Jf"]
    ik = BatchedInvKin(16)
    ik.refreshAndCompute(*[k[n] for n in names])
    for a, n in zip((ik.ddq, ik.dq_cmd, ik.q_step), ("ddq", "dq_cmd", "q_step")):
        np.testing.assert_allclose(a, k[n], rtol=1e-12, atol=1e-12)


def test_wbc_tick_golden():
    from qrw_b200 import BatchedWbcTick

    g = np.load(os.path.join(GOLD, "wbc_tick_b16.npz"))
    t = BatchedWbcTick(16)
    names = ["q", "dq", "f_cmd", "contacts", "pgoals", "vgoals", "agoals"]
    for c in range(2):
        t.compute(*[g[f"{n}{c}"] for n in names])
        assert_forces_close(t.f_with_delta, g[f"f_res{c}"], f"f_with_delta call {c}")
        np.testing.assert_allclose(t.tau_ff, g[f"tau_ff{c}"], rtol=1e-5, atol=1e-4)
        np.testing.assert_allclose(t.qdes, g[f"qdes{c}"], rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(t.vdes, g[f"vdes{c}"], rtol=1e-10, atol=1e-10)
