"""Identities that pin the oracle's Solo12 rigid-body restatement (oracle/solo12_oracle.cpp) in the absence of
Pinocchio: geometry anchors from the reference tree, Jacobian = finite difference of the
placement, rnea linear in the acceleration with a symmetric positive-definite inertia matrix, gravity wrench at rest."""
import numpy as np

MASS = 2.50000279  # src/MPC.cpp:17


def _q0():
    q = np.zeros(19)
    q[6] = 1.0
    return q


def test_geometry_anchors(oracle):
    # straight legs: feet under the shoulders at (+-0.1946, +-0.14695, -0.32) (Controller.py:131-133, test_mpc.py:39-40)
    posf, vf, wf, af, Jf = oracle.solo12_feet(np.zeros((1, 12)), np.zeros((1, 12)))
    p = posf.reshape(3, 4)  # 4 x 3 column-major -> [axis, foot]
    assert np.allclose(p[0], [0.1946, 0.1946, -0.1946, -0.1946])
    assert np.allclose(p[1], [0.14695, -0.14695, 0.14695, -0.14695])
    assert np.allclose(p[2], -0.32)
    # q_init = (0, 0.7, -1.4): base height 2 * 0.16 * cos(0.7) = 0.24474949993103629 (scripts/test_mpc.py:39-40)
    q = np.array([[0.0, 0.7, -1.4] * 2 + [0.0, -0.7, 1.4] * 2])
    p = oracle.solo12_feet(q, np.zeros((1, 12)))[0].reshape(3, 4)
    assert np.allclose(p[2], -0.24474949993103629, atol=1e-12)
    assert np.allclose(np.abs(p[0]), 0.1946, atol=1e-12)


def test_jacobian_is_derivative_of_placement(oracle):
    rng = np.random.default_rng(0)
    q = rng.uniform(-1.0, 1.0, (1, 12))
    Jf = oracle.solo12_feet(q, np.zeros((1, 12)))[4].reshape(12, 12).T  # [row, col]
    h = 1e-6
    for j in range(12):
        dq = np.zeros((1, 12))
        dq[0, j] = h
        pp = oracle.solo12_feet(q + dq, np.zeros((1, 12)))[0].reshape(3, 4)
        pm = oracle.solo12_feet(q - dq, np.zeros((1, 12)))[0].reshape(3, 4)
        num = ((pp - pm) / (2 * h)).T.reshape(12)  # foot-major: rows 3*foot + axis
        assert np.allclose(Jf[:, j], num, atol=1e-8)


def test_foot_velocity_and_acceleration(oracle):
    """vf = Jf dq; the classical foot acceleration af + wf x vf equals the time derivative of Jf dq at constant dq."""
    rng = np.random.default_rng(1)
    q, dq = rng.uniform(-1.0, 1.0, (1, 12)), rng.uniform(-2.0, 2.0, (1, 12))
    posf, vf, wf, af, Jf = oracle.solo12_feet(q, dq)
    J = Jf.reshape(12, 12).T
    v = vf.reshape(3, 4).T.reshape(12)
    assert np.allclose(J @ dq[0], v, atol=1e-12)
    h = 1e-6
    vp = oracle.solo12_feet(q + h * dq, dq)[1].reshape(3, 4).T.reshape(12)
    vm = oracle.solo12_feet(q - h * dq, dq)[1].reshape(3, 4).T.reshape(12)
    acl = (vp - vm) / (2 * h)
    w, a = wf.reshape(3, 4).T, af.reshape(3, 4).T
    assert np.allclose((a + np.cross(w, v.reshape(4, 3))).reshape(12), acl, atol=1e-6)


def test_mass_matrix_properties(oracle):
    M = oracle.solo12_crba(_q0())
    assert np.allclose(M, M.T, atol=1e-12)
    assert np.linalg.eigvalsh(M).min() > 0
    assert np.allclose(np.diag(M)[:3], MASS, atol=1e-8)  # total mass, src/MPC.cpp:17
    rng = np.random.default_rng(2)
    q = _q0()
    q[7:] = rng.uniform(-1, 1, 12)
    quat = rng.normal(size=4)
    q[3:7] = quat / np.linalg.norm(quat)
    M = oracle.solo12_crba(q)
    assert np.allclose(M, M.T, atol=1e-12) and np.linalg.eigvalsh(M).min() > 0
    # rnea is affine in the acceleration: rnea(q, v, a) - rnea(q, v, 0) = M a
    v, a = rng.uniform(-1, 1, 18), rng.uniform(-5, 5, 18)
    t1 = oracle.solo12_rnea(q, v, a)[0]
    t0 = oracle.solo12_rnea(q, v, np.zeros(18))[0]
    assert np.allclose(t1 - t0, M @ a, atol=1e-10)


def test_gravity_wrench_at_rest(oracle):
    """Robot at rest, level base: the base must be pushed up with the weight; tilted base: the same force seen in the base frame."""
    tau = oracle.solo12_rnea(_q0(), np.zeros(18), np.zeros(18))[0]
    assert np.allclose(tau[:3], [0, 0, MASS * 9.81], atol=1e-9)
    q = _q0()
    ang = 0.3
    q[3:7] = [np.sin(ang / 2), 0, 0, np.cos(ang / 2)]  # roll
    tau = oracle.solo12_rnea(q, np.zeros(18), np.zeros(18))[0]
    assert np.allclose(tau[:3], MASS * 9.81 * np.array([0, np.sin(ang), np.cos(ang)]), atol=1e-9)


def test_contact_jacobian_matches_kinematics(oracle):
    """Jc of a foot in contact maps (base twist, joint velocities) to the world velocity of the foot."""
    rng = np.random.default_rng(3)
    q = _q0()
    q[7:] = rng.uniform(-1, 1, 12)
    quat = rng.normal(size=4)
    q[3:7] = quat / np.linalg.norm(quat)
    Jc = oracle.solo12_jc(q, np.ones(4)).reshape(18, 12).T
    x, y, z, w = q[3:7]
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    v = rng.uniform(-1, 1, 18)
    posf, vf, wf, af, Jf = oracle.solo12_feet(q[7:], v[6:])
    p, vrel = posf.reshape(3, 4).T, vf.reshape(3, 4).T
    for foot in range(4):
        want = R @ (v[:3] + np.cross(v[3:6], p[foot]) + vrel[foot])
        assert np.allclose(Jc[3 * foot:3 * foot + 3] @ v, want, atol=1e-12)
    assert np.allclose(oracle.solo12_jc(q, np.array([1.0, 0, 0, 1.0])).reshape(18, 12).T[3:9], 0)
