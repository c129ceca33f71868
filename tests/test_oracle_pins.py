"""Pins of the CPU oracle (oracle/, test infrastructure): what can be checked against reference code or
reference-given numbers in this container, where OSQP/Eigen/Pinocchio are absent.

* the triplet -> CSC layout of ML against the REFERENCE's own src/st_to_cc.cpp (compiled in place into
  oracle/_ref/libst_to_cc.so by oracle/Makefile; skipped when that file is absent);
* the analytic stance forces and the behavioural scenarios of scripts/test_mpc.py:54-190;
* the restated OSQP against the KKT conditions of the QP it claims to have solved;
* the committed golden fixtures (tests/golden, written by tools/make_golden.py).
"""
import os

import numpy as np
import pytest

from conftest import ROOT

GOLD = os.path.join(ROOT, "tests", "golden")
MASS, G = 2.500002861022949, 9.8100004196167
MU = 0.8999999761581421


def ml_triplets(n=16, dt=0.02):
    """Triplets of ML in the order MPC::create_ML emits them (reference src/MPC.cpp:83-146)."""
    r, c, v = [], [], []

    def add(i, j, x):
        r.append(i), c.append(j), v.append(x)

    for k in range(12 * n):
        add(k, k, -1.0)
    for k in range(n - 1):
        for i in range(12):
            add((k + 1) * 12 + i, k * 12 + i, 1.0)
        for j in range(6):
            add((k + 1) * 12 + j, k * 12 + j + 6, dt)
    div = dt / MASS
    for k in range(n):
        for i in range(4):
            for a in range(3):
                add(12 * k + 6 + a, 12 * (n + k) + a + 3 * i, div)
        for i in range(12):
            for a in range(3):
                add(12 * k + 9 + a, 12 * (n + k) + i, 8.0)
    for i in range(12 * n, 24 * n):
        add(i, i, 1.0)
    off = 24 * n
    a_ = [0, 1, 2, 3, 0, 1, 2, 3, 4]
    b_ = [0, 0, 1, 1, 2, 2, 2, 2, 2]
    c_ = [1.0, -1.0, 1.0, -1.0, -MU, -MU, -MU, -MU, -1.0]
    for k in range(n):
        for i in range(4):
            for j in range(9):
                add(off + 20 * k + 5 * i + a_[j], 12 * (n + k) + 3 * i + b_[j], c_[j])
    return np.array(r, np.int32), np.array(c, np.int32), np.array(v)


def test_ml_layout_matches_reference_st_to_cc(oracle):
    """the CSC pattern the oracle (and the kernels' index maps) assume is the one the
    reference's converter produces from the reference's triplets: 1998 nnz, [7,7,10] per force foot."""
    if not oracle.ref_st_to_cc_available():
        pytest.skip("oracle/_ref/libst_to_cc.so not built (needs /root/reference at build time)")
    from qrw_b200 import workloads as wl

    r, c, v = ml_triplets()
    icc, ccc, acc = oracle.ref_st_to_cc(r, c, v, 384)
    assert icc.size == 1998 and ccc[-1] == 1998
    o = oracle.OracleMPC(1)
    xr, fs = wl.config1_single_trot()
    o.run(0, xr, fs)
    x, ri, cp = o.ml(0)
    assert np.array_equal(ri, icc) and np.array_equal(cp, ccc)
    # values: everything but the 16 x 36 angular B entries (placeholder 8.0 in the triplets) and the S flags
    same = (acc != 8.0) & ~((ri >= 192) & (ri < 384))
    assert np.array_equal(x[same], acc[same])
    counts = np.diff(ccc)[192:].reshape(16, 4, 3)
    assert (counts == np.array([7, 7, 10])).all()


def test_four_stance_analytic_force(oracle):
    """scripts/test_mpc.py:54-85 (four feet down, robot at rest): equal forces, fz = m g / 4 = 6.131257 N,
    predicted next state = reference state."""
    from qrw_b200 import workloads as wl

    o = oracle.OracleMPC(1)
    xr, fs = wl.four_stance()
    for k in range(3):
        o.run(k, xr, fs)
    R = o.get_latest_result().reshape(16, 24)
    f = R[0, 12:].reshape(4, 3)
    assert np.allclose(f, f[0], atol=1e-4)
    # the exact-value assertion is commented out in the reference (test_mpc.py:85): the horizon cost bends fz by ~1e-2
    assert abs(f[0, 2] - MASS * G / 4) < 2e-2 and abs(MASS * G / 4 - 6.131257) < 1e-6
    assert np.allclose(R[0, :12], xr.reshape(17, 12)[1], atol=1e-3)


def test_four_stance_offset_converges(oracle):
    """scripts/test_mpc.py:87-110: start 2 cm low with a roll offset, apply the predicted state for 100 calls."""
    from qrw_b200 import workloads as wl

    o = oracle.OracleMPC(1)
    x0 = np.zeros(12)
    x0[2], x0[3] = wl.Z_INIT - 0.02, 0.05
    for k in range(100):
        xr, fs = wl.four_stance(x0)
        o.run(k, xr, fs)
        x0 = o.get_latest_result()[0, :12].copy()
    assert np.allclose(x0, wl.four_stance()[0].reshape(17, 12)[1], atol=1e-3)


def test_trot_stance_force(oracle):
    """Trot at rest: the two stance feet share the weight, fz = m g / 2 = 12.262515 N."""
    from qrw_b200 import workloads as wl

    o = oracle.OracleMPC(1)
    xr, fs = wl.config1_single_trot()
    o.run(0, xr, fs)
    o.run(1, xr, fs)
    f = o.get_latest_result().reshape(16, 24)[:, 12:].reshape(16, 4, 3)
    assert abs(f[0, 0, 2] + f[0, 3, 2] - MASS * G) < 0.2
    assert abs(MASS * G / 2 - 12.262515) < 1e-5
    assert np.abs(f[0, 1:3]).max() < 1e-4  # swing feet (FR, HL) pinned to zero by the S rows
    assert np.abs(f[8, [0, 3]]).max() < 1e-4 and f[8, 1, 2] > 5.0


def test_restated_osqp_satisfies_kkt(oracle):
    """The ADMM restatement on random strictly convex QPs: the returned (x, y) meets the KKT conditions."""
    rng = np.random.default_rng(0)
    for trial in range(5):
        n, m = 8, 12
        L = rng.normal(size=(n, n))
        P = L @ L.T + 0.1 * np.eye(n)
        A = rng.normal(size=(m, n))
        q = rng.normal(size=n)
        l = -rng.uniform(0.1, 1.0, m)
        u = rng.uniform(0.1, 1.0, m)
        l[:3] = u[:3] = 0.2  # equality rows
        x, y, it, st = oracle.osqp_dense(P, A, q, l, u, eps=1e-9)
        assert st == 1
        z = A @ x
        assert (z >= l - 1e-6).all() and (z <= u + 1e-6).all()
        assert np.abs(P @ x + q + A.T @ y).max() < 1e-6
        # complementarity: y > 0 only on the upper bound, y < 0 only on the lower bound
        assert np.abs(np.maximum(y, 0) * (u - z)).max() < 1e-5
        assert np.abs(np.minimum(y, 0) * (z - l)).max() < 1e-5


@pytest.mark.parametrize("name,B", [("mpc_config1_single_trot", 1), ("mpc_trot_b8", 8), ("mpc_mixed_b10", 10)])
def test_oracle_reproduces_mpc_golden(oracle, name, B):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    o = oracle.OracleMPC(B)
    for t in range(int(g["ticks"])):
        o.run(t, g[f"xref{t}"], g[f"fsteps{t}"], 8)
        assert np.array_equal(o.get_gait(), g[f"gait{t}"]) and np.array_equal(o.get_Sgait(), g[f"sgait{t}"])
        assert np.array_equal(o.info()["iters"], g[f"iters{t}"])
        np.testing.assert_allclose(o.get_latest_result(), g[f"result{t}"], rtol=0, atol=1e-9)


def test_oracle_reproduces_wbc_invkin_golden(oracle):
    g = np.load(os.path.join(GOLD, "wbc_b16.npz"))
    o = oracle.OracleWBC(16)
    for c in range(3):
        o.run(g[f"M{c}"], g[f"Jc{c}"], g[f"f_cmd{c}"], g[f"RNEA{c}"], 4)
        f, d, H = o.results()
        assert np.array_equal(o.info()["iters"], g[f"iters{c}"])
        np.testing.assert_allclose(f, g[f"f_res{c}"], rtol=0, atol=1e-10)
        np.testing.assert_allclose(d, g[f"ddq_res{c}"], rtol=0, atol=1e-9)
    k = np.load(os.path.join(GOLD, "invkin_b16.npz"))
    names = ["contacts", "goals", "vgoals", "agoals", "posf", "vf", "wf", "af", "Jf"]
    outs = oracle.invkin(*[k[n] for n in names])
    for a, n in zip(outs, ("ddq", "dq_cmd", "q_step")):
        np.testing.assert_allclose(a, k[n], rtol=1e-13, atol=1e-13)


def test_oracle_reproduces_wbc_tick_golden(oracle):
    g = np.load(os.path.join(GOLD, "wbc_tick_b16.npz"))
    o = oracle.OracleWBC(16)
    names = ["q", "dq", "f_cmd", "contacts", "pgoals", "vgoals", "agoals"]
    for c in range(2):
        r = oracle.wbc_tick(o, *[g[f"{n}{c}"] for n in names], nthreads=4)
        for k in ("tau_ff", "qdes", "vdes", "f_res"):
            np.testing.assert_allclose(r[k], g[f"{k}{c}"], rtol=0, atol=1e-9)


# ---------------------------------------------------------------------------------------------------------------------
# The one OSQP output this image can be compared with: the run printed in OSQP's own documentation (osqp.org, "Examples ->
# Setup and solve", OSQP v0.6.0; Python: prob.setup(P, q, A, l, u, alpha=1.0); prob.solve()):
#
#     iter   objective    pri res    dua res    rho        time
#        1  -4.9384e-03   1.00e+00   2.00e+02   1.00e-01   ...
#       50   1.8800e+00   1.91e-07   7.50e-07   1.38e+00   ...
#     status: solved   number of iterations: 50   optimal objective: 1.8800   optimal rho estimate: 1.36e+00
#
# No network and no OSQP exist here: the figures are written down from the builder's memory of that page.  They were NOT fitted:
# the restatement reproduces every printed digit of all eleven of them, which a faulty recollection or a faulty restatement
# would not.  That run exercises the Ruiz equilibration and cost scaling (iteration 1 depends on D, E, c), rho of an equality
# row (x 1000), the KKT solve, the unscaled residuals, the termination test (not met at iteration 25, met at 50), the rho
# estimate and update (0.1 -> 1.38 at iteration 25; in that run OSQP's timing-based rule chose the interval, here it is set).
# ---------------------------------------------------------------------------------------------------------------------
OSQP_DOC_P = np.array([[4.0, 1.0], [1.0, 2.0]])
OSQP_DOC_Q = np.array([1.0, 1.0])
OSQP_DOC_A = np.array([[1.0, 1.0], [1.0, 0.0], [0.0, 1.0]])
OSQP_DOC_L = np.array([1.0, 0.0, 0.0])
OSQP_DOC_U = np.array([1.0, 0.7, 0.7])


def _osqp_line(it, x, info):
    obj = 0.5 * x @ OSQP_DOC_P @ x + OSQP_DOC_Q @ x
    return "%4d %12.4e %10.2e %10.2e %10.2e" % (it, obj, info["pri_res"], info["dua_res"], info["rho"])


def test_restated_osqp_reproduces_the_run_printed_in_osqps_documentation(oracle):
    args = (OSQP_DOC_P, OSQP_DOC_A, OSQP_DOC_Q, OSQP_DOC_L, OSQP_DOC_U)
    x, y, it, st, info = oracle.osqp_dense_ex(*args, alpha=1.0, max_iter=1)
    assert _osqp_line(it, x, info).split() == ["1", "-4.9384e-03", "1.00e+00", "2.00e+02", "1.00e-01"]
    x, y, it, st, info = oracle.osqp_dense_ex(*args, alpha=1.0, adaptive_rho_interval=25)
    assert st == 1  # OSQP_SOLVED
    assert _osqp_line(it, x, info).split() == ["50", "1.8800e+00", "1.91e-07", "7.50e-07", "1.38e+00"]
    assert info["rho_updates"] == 1 and "%.2e" % info["rho_estimate"] == "1.36e+00"
    np.testing.assert_allclose(x, [0.3, 0.7], atol=1e-6)         # the solution the page prints
    np.testing.assert_allclose(y, [-2.9, 0.0, 0.2], atol=1e-6)
    # the interval is the one free choice and it is not a fit: with any other multiple of 25 the run ends differently
    for interval in (50, 100):
        _, _, it2, _, info2 = oracle.osqp_dense_ex(*args, alpha=1.0, adaptive_rho_interval=interval)
        assert (it2, "%.2e" % info2["rho"]) != (50, "1.38e+00")
