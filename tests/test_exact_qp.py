"""An independent EXACT solve of the QP that MPC::run hands to OSQP (reference src/MPC.cpp:74-391: n = 384 variables,
m = 704 rows), to bound |restated ADMM - optimum| on the real MPC problem.

OSQP itself is absent from this container, so the oracle's ADMM (oracle/osqp_restated.hpp) cannot be diffed against the
solver the reference links.  What CAN be checked without it:
  1. the QP data the oracle assembles (ML in CSC order, bounds) equals an assembly written here from the reference's
     formulas in plain numpy (dynamics, B_k = dt I^-1 skew(r - c), cone rows) -- a second, independent restatement;
  2. the oracle's returned forces are within a stated distance of the TRUE optimum of that QP, computed by a
     primal-dual interior-point method with an active-set polish (KKT residual <= 1e-9), which shares no code with
     the ADMM restatement.
eps_abs = eps_rel = 1e-6 (MPC.cpp:527-528) stops the ADMM before the weakly determined directions (force regulariser
5e-5) have converged, so the distance is not round-off: the measured bound is asserted with a margin and quoted in
BASELINE.md.
"""
import numpy as np
import pytest

MASS, MU = 2.500002861022949, 0.8999999761581421
G8 = 9.8100004196167
GI = np.array([[3.09249e-2, -8.00101e-7, 1.865287e-5], [-8.00101e-7, 5.106100e-2, 1.245813e-4],
               [1.865287e-5, 1.245813e-4, 6.939757e-2]])
W = np.array([2.0, 2.0, 20.0, 0.25, 0.25, 10.0, 0.20000000298023224, 0.20000000298023224, 0.20000000298023224, 0.0, 0.0,
              0.30000001192092896])
WF = 4.999999873689376e-05
N, DT = 16, 0.02


def skew(v):
    return np.array([[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]])


def assemble(xref, fsteps, sgait):
    """Dense A (704 x 384), l, u of the update path (MPC.cpp:418-496) from xref [12, 17], fsteps [20, 12] and the S flags."""
    n = N
    A = np.zeros((44 * n, 24 * n))
    Ad = np.eye(12)
    Ad[:6, 6:] = DT * np.eye(6)
    for k in range(n):
        A[12 * k:12 * k + 12, 12 * k:12 * k + 12] = -np.eye(12)
        if k + 1 < n:
            A[12 * (k + 1):12 * (k + 1) + 12, 12 * k:12 * k + 12] = Ad
        yaw = xref[5, k]
        c, s = np.cos(yaw), np.sin(yaw)
        R = np.array([[c, -s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])
        Iinv = np.linalg.inv(R.T @ GI @ R)
        com = xref[0:3, k] + np.array([0.0, 0.0, -0.03])
        Bk = np.zeros((12, 12))
        for f in range(4):
            Bk[6:9, 3 * f:3 * f + 3] = (DT / MASS) * np.eye(3)
            Bk[9:12, 3 * f:3 * f + 3] = DT * Iinv @ skew(fsteps[k, 3 * f:3 * f + 3] - com)
        A[12 * k:12 * k + 12, 12 * (n + k):12 * (n + k) + 12] = Bk
    for j in range(12 * n):
        A[12 * n + j, 12 * n + j] = sgait[j]
    Cf = np.array([[1, 0, -MU], [-1, 0, -MU], [0, 1, -MU], [0, -1, -MU], [0, 0, -1.0]])
    for k in range(n):
        for f in range(4):
            r0, c0 = 24 * n + 20 * k + 5 * f, 12 * (n + k) + 3 * f
            A[r0:r0 + 5, c0:c0 + 3] = Cf
    g = np.zeros(12)
    g[8] = -G8 * DT
    l = np.zeros(44 * n)
    u = np.zeros(44 * n)
    for k in range(n):
        prev = xref[:, k]
        rhs = xref[:, k + 1] - Ad @ prev - g if k > 0 else -Ad @ xref[:, 0] - g + xref[:, 1]
        l[12 * k:12 * k + 12] = u[12 * k:12 * k + 12] = rhs
    l[24 * n:] = -np.inf
    l[24 * n + 4::5] = -25.0
    return A, l, u


def exact_qp(Pd, A, l, u, tol=1e-11):
    """min 1/2 x' diag(Pd) x  s.t.  l <= A x <= u, by a primal-dual interior point method (Mehrotra predictor-corrector,
    dense KKT solves) on {A_eq x = b, G x <= h}.  Returns x, the largest KKT residual (stationarity, primal feasibility,
    complementarity gap) and the number of active inequality rows.  The swing-foot forces are pinned by the S rows AND sit
    on their cone rows, so the active set is degenerate: no active-set polish, the interior point iterates are taken to 1e-11."""
    n_full = A.shape[1]
    # presolve: an equality row with one entry pins its variable (the S rows pin the swing-foot forces to 0); the cone rows
    # of those feet become 0 <= 0 and leave no strict interior, so pinned variables and constant rows are removed first
    eq = np.isfinite(l) & (u - l < 1e-12)
    single = eq & ((A != 0).sum(1) == 1)
    pinned = np.zeros(n_full, bool)
    xfix = np.zeros(n_full)
    for r in np.where(single)[0]:
        c = int(np.flatnonzero(A[r])[0])
        pinned[c], xfix[c] = True, u[r] / A[r, c]
    shift = A @ xfix
    l, u = l - shift, u - shift
    keepc = ~pinned
    A = A[:, keepc]
    Pd = Pd[keepc]
    const = (A != 0).sum(1) == 0
    assert (l[const] <= 1e-12).all() and (u[const] >= -1e-12).all(), "presolve: a constant row is infeasible"
    A, l, u = A[~const], l[~const], u[~const]
    n = A.shape[1]
    eq = np.isfinite(l) & (u - l < 1e-12)
    ine = ~eq
    Ae, be = A[eq], u[eq]
    rows = np.where(ine)[0]
    G = [A[rows[np.isfinite(u[rows])]]]
    h = [u[rows[np.isfinite(u[rows])]]]
    lo = rows[np.isfinite(l[rows])]
    G.append(-A[lo])
    h.append(-l[lo])
    G, h = np.vstack(G), np.concatenate(h)
    me, mi = Ae.shape[0], G.shape[0]
    P = np.diag(Pd)
    x = np.zeros(n)
    s = np.ones(mi)
    z = np.ones(mi)
    y = np.zeros(me)
    reg = 1e-11
    for it in range(200):
        rd = P @ x + Ae.T @ y + G.T @ z
        rp = Ae @ x - be
        rc = G @ x + s - h
        mu = s @ z / mi
        if max(np.abs(rd).max(), np.abs(rp).max(), np.abs(rc).max()) < tol and mu < tol:
            break
        Wd = z / s
        K = np.block([[P + G.T @ (Wd[:, None] * G) + reg * np.eye(n), Ae.T], [Ae, -reg * np.eye(me)]])

        def solve(rd_, rp_, rc_, rs_):
            # rs_ = residual of s z = target (elementwise)
            rhs1 = -rd_ - G.T @ ((-rs_ + z * rc_) / s)
            sol = np.linalg.solve(K, np.concatenate([rhs1, -rp_]))
            dx, dy = sol[:n], sol[n:]
            ds = -rc_ - G @ dx
            dz = (-rs_ - z * ds) / s
            return dx, dy, ds, dz

        dx, dy, ds, dz = solve(rd, rp, rc, s * z)

        def step(v, dv):
            neg = dv < 0
            return min(1.0, 0.995 * np.min(-v[neg] / dv[neg])) if neg.any() else 1.0

        a = min(step(s, ds), step(z, dz))
        mu_aff = (s + a * ds) @ (z + a * dz) / mi
        sig = (mu_aff / mu) ** 3
        dx, dy, ds, dz = solve(rd, rp, rc, s * z + ds * dz - sig * mu)
        a = min(step(s, ds), step(z, dz))
        x, y, s, z = x + a * dx, y + a * dy, s + a * ds, z + a * dz
    res = max(np.abs(P @ x + Ae.T @ y + G.T @ z).max(), np.abs(Ae @ x - be).max(), np.maximum(G @ x - h, 0.0).max(), s @ z / mi)
    xf = xfix.copy()
    xf[keepc] = x
    return xf, res, int((z > s).sum())


def _qp_of(o, b):
    x, ri, cp = o.ml(b)
    A = np.zeros((704, 384))
    for c in range(384):
        A[ri[cp[c]:cp[c + 1]], c] = x[cp[c]:cp[c + 1]]
    l, u = o.bounds(b)
    l = np.where(l < -1e29, -np.inf, l)
    return A, l, u


def test_oracle_qp_data_equal_an_independent_assembly(oracle):
    """Update path: ML values and bounds of the oracle = the numpy assembly above, for trot robots (16 active rows)."""
    from qrw_b200 import workloads as wl

    B = 6
    w = wl.MpcWorkload(B, 11, gaits=("trot", "walk", "pacing", "static"), ticks=3)
    o = oracle.OracleMPC(B)
    xr, fs = w.inputs(0)
    o.run(0, xr, fs, 4)
    x0 = o.get_latest_result()[:, :12].copy()
    xr, fs = w.inputs(1, x0)
    o.run(1, xr, fs, 4)
    sg = o.get_Sgait()
    for b in range(B):
        A, l, u = _qp_of(o, b)
        A2, l2, u2 = assemble(xr[b].reshape(17, 12).T, fs[b].reshape(12, 20).T, sg[b])
        np.testing.assert_allclose(A, A2, rtol=1e-12, atol=1e-15)
        np.testing.assert_allclose(u, u2, rtol=0, atol=1e-14)
        assert np.array_equal(np.isfinite(l), np.isfinite(l2))
        np.testing.assert_allclose(l[np.isfinite(l)], l2[np.isfinite(l2)], rtol=0, atol=1e-14)


# measured here: max |f_oracle - f_exact| over all 192 forces of a solve = 1.3e-2 N (config 1 cold), 5.9e-3 N (config 1
# warm), 2.4e-2 N (config-3 sample, first step 2.2e-2 N).  The ADMM stops at eps_abs = eps_rel = 1e-6 while the force
# regulariser is 5e-5: a force error df moves the dual residual by only 5e-5 df, so the stopping test admits df of the
# order of 1e-2 N in the directions the dynamics rows do not see.  That is a property of the reference's settings
# (MPC.cpp:527-528), not of the restatement: with eps = 1e-10 the same ADMM lands on the exact optimum (test below).
# The bound is therefore how far "OSQP at the reference's eps" may sit from the optimum, and why the GPU parity tests
# compare iterate sequences (iteration counts, 1e-4 N on forces) rather than optima.
BOUND_ALL, BOUND_FIRST = 5e-2, 5e-2


def _check(o, b, res):
    A, l, u = _qp_of(o, b)
    Pd = np.concatenate([np.tile(W, N), np.full(192, WF)])
    x, kkt, nact = exact_qp(Pd, A, l, u)
    assert kkt < 1e-9, f"exact solver: KKT residual {kkt:.2e}"
    f_exact = x[192:].reshape(16, 12)
    f_orc = res.reshape(16, 24)[:, 12:]
    return np.abs(f_orc - f_exact).max(), np.abs(f_orc[0] - f_exact[0]).max()


def test_oracle_admm_is_within_a_bound_of_the_exact_optimum_config1(oracle):
    from qrw_b200 import workloads as wl

    o = oracle.OracleMPC(1)
    xr, fs = wl.config1_single_trot()
    for k in range(2):
        o.run(k, xr, fs)
        e_all, e_first = _check(o, 0, o.get_latest_result()[0])
        print(f"config 1 call {k}: |oracle - exact| all forces {e_all:.3e} N, first step {e_first:.3e} N")
        assert e_all < BOUND_ALL and e_first < BOUND_FIRST


def test_oracle_admm_is_within_a_bound_of_the_exact_optimum_config3_sample(oracle):
    from qrw_b200 import workloads as wl

    B = 10
    w = wl.MpcWorkload(B, 2, gaits=wl.GAITS, ticks=3)
    o = oracle.OracleMPC(B)
    x0 = None
    worst = [0.0, 0.0]
    for t in range(2):
        xr, fs = w.inputs(t, x0)
        o.run(t, xr, fs, 8)
        res = o.get_latest_result()
        for b in range(B):
            e_all, e_first = _check(o, b, res[b])
            worst = [max(worst[0], e_all), max(worst[1], e_first)]
        x0 = res[:, :12].copy()
    print(f"config-3 sample: |oracle - exact| all forces {worst[0]:.3e} N, first step {worst[1]:.3e} N")
    assert worst[0] < BOUND_ALL and worst[1] < BOUND_FIRST


def test_restated_admm_converges_to_the_exact_optimum_of_the_mpc_qp(oracle):
    """The restated OSQP iteration (scaling, rho rules, relaxation) run to eps = 1e-10 on the MPC QP of config 1 reaches
    the interior-point optimum to 1e-5 N (measured 2e-6): its fixed point is the QP's solution, whatever path it takes."""
    from qrw_b200 import workloads as wl

    o = oracle.OracleMPC(1)
    xr, fs = wl.config1_single_trot()
    o.run(0, xr, fs)
    o.run(1, xr, fs)
    A, l, u = _qp_of(o, 0)
    Pd = np.concatenate([np.tile(W, N), np.full(192, WF)])
    x_exact, kkt, _ = exact_qp(Pd, A, l, u)
    assert kkt < 1e-9
    x, y, it, st = oracle.osqp_dense(np.diag(Pd), A, np.zeros(384), np.where(np.isfinite(l), l, -1e30), u, eps=1e-10, max_iter=200000)
    assert st == 1
    err = np.abs(x[192:] - x_exact[192:]).max()
    print(f"ADMM at eps 1e-10 vs interior point: {err:.3e} N after {it} iterations")
    assert err < 1e-5
