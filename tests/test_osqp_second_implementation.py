"""A SECOND implementation of the OSQP v0.6 iteration, written here in dense numpy from the published algorithm (Stellato et al.
2020) and OSQP's scaling / termination / rho rules, against the oracle's C++ restatement (oracle/osqp_restated.hpp) on small random
QPs: same number of ADMM iterations, same solution.  It shares no code with the oracle (dense LU instead of sparse LDL', numpy
reductions instead of CSC loops), so a transcription slip in either -- column norms of the upper-triangular P, the order of the
Ruiz and cost-scaling steps, rho of equality / loose rows, unscaled residuals, the rho estimate -- shows up as a different iteration
count.  No OSQP binary exists in this image: this is a consistency check of the restatement, not a pin against OSQP."""
import numpy as np
import pytest

INF = 1e30  # OSQP_INFTY
MIN_S, MAX_S = 1e-4, 1e4


def _limit(v):
    v = np.where(v < MIN_S, 1.0, v)
    return np.where(v > MAX_S, MAX_S, v)


def osqp_numpy(P, A, q, l, u, eps, max_iter=4000, rho=0.1, sigma=1e-6, alpha=1.6, check=25, interval=200, tol=5.0):
    P, A, q, l, u = (np.array(a, dtype=np.float64) for a in (P, A, q, l, u))
    m, n = A.shape
    D, E, c = np.ones(n), np.ones(m), 1.0
    for _ in range(10):  # scale_data: Ruiz on the KKT matrix, then the cost normalisation
        dt = 1.0 / np.sqrt(_limit(np.maximum(np.abs(P).max(0), np.abs(A).max(0))))
        et = 1.0 / np.sqrt(_limit(np.abs(A).max(1)))
        P = (dt[:, None] * P) * dt[None, :]
        A = (et[:, None] * A) * dt[None, :]
        q = dt * q
        D, E = D * dt, E * et
        cn = sum(np.abs(P).max(0)) / n
        ct = 1.0 / float(_limit(np.array(max(cn, float(_limit(np.array(np.abs(q).max())))))))
        P, q, c = P * ct, q * ct, c * ct
    l = np.where(l <= -INF, -INF, E * l)
    u = np.where(u >= INF, INF, E * u)
    loose = (l < -INF * MIN_S) & (u > INF * MIN_S)
    eq = (u - l) < 1e-4

    def rho_vec(r):
        return np.where(loose, 1e-6, np.where(eq, 1e3 * r, r))

    x, z, y = np.zeros(n), np.zeros(m), np.zeros(m)
    noisy = False
    rv = rho_vec(rho)
    K = np.block([[P + sigma * np.eye(n), A.T], [A, -np.diag(1.0 / rv)]])
    for it in range(1, max_iter + 1):
        xp, zp = x, z
        sol = np.linalg.solve(K, np.concatenate([sigma * xp - q, zp - y / rv]))
        xt, zt = sol[:n], zp + (sol[n:] - y) / rv
        x = alpha * xt + (1 - alpha) * xp
        zr = alpha * zt + (1 - alpha) * zp
        z = np.clip(zr + y / rv, l, u)
        y = y + rv * (zr - z)
        if it % check == 0 or it % interval == 0:
            Ax, Px, Aty = A @ x, P @ x, A.T @ y
            if it % check == 0:
                pri = np.abs((Ax - z) / E).max()
                dua = np.abs((Px + q + Aty) / D).max() / c
                eps_p = eps + eps * max(np.abs(z / E).max(), np.abs(Ax / E).max())
                eps_d = eps + eps * max(np.abs(Px / D).max(), np.abs(Aty / D).max(), np.abs(q / D).max()) / c
                if pri < eps_p and dua < eps_d:
                    return D * x, E * y / c, it, noisy
            if it % interval == 0:  # compute_rho_estimate on the SCALED quantities
                pr = np.abs(Ax - z).max() / (max(np.abs(z).max(), np.abs(Ax).max()) + 1e-10)
                dr = np.abs(Px + q + Aty).max() / (max(np.abs(q).max(), np.abs(Aty).max(), np.abs(Px).max()) + 1e-10)
                noisy = noisy or min(pr, dr) < 1e-8  # an estimate taken from a residual at round-off level is not reproducible
                new = min(max(rho * np.sqrt(pr / (dr + 1e-10)), 1e-6), 1e6)
                if new > rho * tol or new < rho / tol:
                    rho, rv = new, rho_vec(new)
                    K = np.block([[P + sigma * np.eye(n), A.T], [A, -np.diag(1.0 / rv)]])
    return D * x, E * y / c, max_iter, noisy


def _random_qp(rng, n, m):
    G = rng.normal(size=(n, n))
    P = G @ G.T * rng.uniform(0.01, 10.0) + np.diag(rng.uniform(1e-3, 1.0, n))
    P[np.abs(P) < 0.3 * np.abs(P).mean()] = 0.0  # some structural zeros; keep it symmetric
    P = np.triu(P) + np.triu(P, 1).T
    P += (abs(min(np.linalg.eigvalsh(P).min(), 0.0)) + 1e-2) * np.eye(n)
    A = rng.normal(size=(m, n)) * (rng.uniform(size=(m, n)) < 0.6) * 10.0 ** rng.uniform(-2, 2, (m, 1))
    A[np.abs(A).sum(1) == 0, 0] = 1.0
    q = rng.normal(size=n) * 10.0 ** rng.uniform(-1, 2)
    x0 = rng.normal(size=n)
    l, u = A @ x0 - rng.uniform(0.1, 2.0, m), A @ x0 + rng.uniform(0.1, 2.0, m)
    kind = rng.integers(0, 5, m)
    l = np.where(kind == 0, A @ x0, l)  # equality rows (rho x 1000), satisfied by x0 like every other row
    u = np.where(kind == 0, A @ x0, u)
    u = np.where(kind == 1, INF, u)    # one-sided rows
    l = np.where(kind == 2, -INF, l)
    return P, A, q, l, u


@pytest.mark.parametrize("seed", range(24))
def test_second_implementation_follows_the_same_iterates(oracle, seed):
    rng = np.random.default_rng(100 + seed)
    n, m = int(rng.integers(3, 12)), int(rng.integers(2, 20))
    P, A, q, l, u = _random_qp(rng, n, m)
    eps = 1e-6
    x2, y2, it2, noisy = osqp_numpy(P, A, q, l, u, eps)
    x1, y1, it1, st = oracle.osqp_dense(P, A, q, l, u, eps=eps, max_iter=4000)
    assert st == 1, f"the QP is feasible by construction; restated OSQP returned status {st}"
    if noisy:
        # a rho update was estimated from a residual at round-off level (1e-11 against 1e-10 of regularisation): dense LU and
        # sparse LDL' legitimately take different paths from there; both must still arrive at the same optimum
        assert it2 < 4000
        np.testing.assert_allclose(x1, x2, rtol=1e-3, atol=1e-4)
        return
    assert it1 == it2, f"ADMM iterations: restatement {it1}, second implementation {it2}"
    np.testing.assert_allclose(x1, x2, rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(y1, y2, rtol=1e-6, atol=1e-8)


def test_second_implementation_reproduces_the_run_printed_in_osqps_documentation():
    """The same published run as tests/test_oracle_pins.py::test_restated_osqp_reproduces_the_run_printed_in_osqps_documentation
    (P = [[4, 1], [1, 2]], q = [1, 1], A = [[1, 1], [1, 0], [0, 1]], l = [1, 0, 0], u = [1, 0.7, 0.7], alpha = 1.0), through
    the numpy implementation: objective after iteration 1, and 50 iterations to the printed solution."""
    P, q = np.array([[4.0, 1.0], [1.0, 2.0]]), np.array([1.0, 1.0])
    A, l, u = np.array([[1.0, 1.0], [1.0, 0.0], [0.0, 1.0]]), np.array([1.0, 0.0, 0.0]), np.array([1.0, 0.7, 0.7])
    x, _, _, _ = osqp_numpy(P, A, q, l, u, 1e-3, max_iter=1, alpha=1.0, check=1000, interval=1000)
    assert "%.4e" % (0.5 * x @ P @ x + q @ x) == "-4.9384e-03"
    x, y, it, _ = osqp_numpy(P, A, q, l, u, 1e-3, alpha=1.0, interval=25)
    assert it == 50 and "%.4e" % (0.5 * x @ P @ x + q @ x) == "1.8800e+00"
    np.testing.assert_allclose(x, [0.3, 0.7], atol=1e-6)
    np.testing.assert_allclose(y, [-2.9, 0.0, 0.2], atol=1e-6)
