"""An INDEPENDENT pin of the Solo12 rigid-body constants (include/solo12_model.h), which kernel and oracle share (VERDICT r1
item 2d / SURVEY.md row a18).  The reference loads the model through Pinocchio from a URDF that is not in its tree, but the
tree does hold numbers DERIVED from that model.  This test rebuilds them from the header with a few lines of numpy written
here (no kernel code, no oracle code) and compares:

  total mass 2.50000279                      src/MPC.cpp:17, scripts/crocoddyl_class/MPC_crocoddyl.py:29
  base-link inertia diag                      scripts/crocoddyl_class/MPC_crocoddyl.py:32 (commented line)
  shoulders (+-0.1946, +-0.14695)             scripts/Controller.py:131-133, src/InvKin.cpp:9
  leg segments: stance height 0.2447495 at q_init = (0, 0.7, -1.4)     scripts/test_mpc.py:39-40
  composite inertia of the robot about its centre of mass (the MPC's gI)  src/MPC.cpp:25-26, MPC_crocoddyl.py:33-35
  height of the centre of mass under the base (the MPC's offset_CoM = -0.03 "approximation")   src/MPC.cpp:21

The last two are the strong ones: they depend on every link mass, centre of mass and inertia in the header, and the header's
values were NOT fitted to them (only the lower-leg mass was fitted, to the total mass).  They agree to 4 % / 5 mm, the
level at which the reference's own rounded, configuration-dependent figures can be reproduced; a wrong link mass, a swapped
axis or a mirrored offset moves them by far more.  The oracle's crba is then checked against the same numpy rebuild, so the
oracle's algorithm is pinned against a second implementation as well."""
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GI_REF = np.array([[3.09249e-2, -8.00101e-7, 1.865287e-5], [-8.00101e-7, 5.106100e-2, 1.245813e-4],
                   [1.865287e-5, 1.245813e-4, 6.939757e-2]])  # src/MPC.cpp:25-26
Q_INIT = np.array([0.0, 0.7, -1.4] * 2 + [0.0, -0.7, 1.4] * 2)  # scripts/main_solo12_control.py


def _model():
    txt = open(os.path.join(ROOT, "include", "solo12_model.h")).read()
    return {m.group(1): float(m.group(2)) for m in re.finditer(r"#define\s+S12_(\w+)\s+\(?(-?[0-9.eE+-]+)\)?", txt)}


def _rx(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]])


def _ry(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def _links(C, q12):
    """(mass, centre of mass, inertia about it) of the 13 bodies in the base frame, and the four foot points."""
    bodies = [(C["BASE_MASS"], np.zeros(3), np.diag([C["BASE_IXX"], C["BASE_IYY"], C["BASE_IZZ"]]))]
    feet = []
    for leg in range(4):
        sx, sy = (1.0 if leg < 2 else -1.0), (-1.0 if leg & 1 else 1.0)
        q = q12[3 * leg:3 * leg + 3]
        R0 = _rx(q[0])
        R1 = R0 @ _ry(q[1])
        R2 = R1 @ _ry(q[2])
        o0 = np.array([sx * C["HAA_X"], sy * C["HAA_Y"], 0.0])
        o1 = o0 + R0 @ np.array([0.0, sy * C["HFE_Y"], 0.0])
        o2 = o1 + R1 @ np.array([0.0, sy * C["KFE_Y"], C["KFE_Z"]])
        feet.append(o2 + R2 @ np.array([0.0, sy * C["FOOT_Y"], C["FOOT_Z"]]))

        def inertia(p, sxy, sxz, syz):
            return np.array([[C[p + "IXX"], sxy * C[p + "IXY"], sxz * C[p + "IXZ"]], [sxy * C[p + "IXY"], C[p + "IYY"], syz * C[p + "IYZ"]],
                             [sxz * C[p + "IXZ"], syz * C[p + "IYZ"], C[p + "IZZ"]]])

        for R, o, p, com, I in ((R0, o0, "L0_", [sx * C["L0_CX"], sy * C["L0_CY"], C["L0_CZ"]], inertia("L0_", sx * sy, sx, sy)),
                                (R1, o1, "L1_", [C["L1_CX"], sy * C["L1_CY"], C["L1_CZ"]], inertia("L1_", sy, 1.0, sy)),
                                (R2, o2, "L2_", [C["L2_CX"], sy * C["L2_CY"], C["L2_CZ"]], inertia("L2_", sy, 1.0, sy))):
            bodies.append((C[p + "MASS"], o + R @ np.array(com), R @ I @ R.T))
    return bodies, np.array(feet)


def _composite(bodies):
    m = sum(b[0] for b in bodies)
    c = sum(b[0] * b[1] for b in bodies) / m
    I = np.zeros((3, 3))
    for mb, cb, Ib in bodies:
        d = cb - c
        I += Ib + mb * (d @ d * np.eye(3) - np.outer(d, d))
    return m, c, I


def test_mass_base_inertia_and_geometry_against_the_reference_tree():
    C = _model()
    bodies, feet = _links(C, np.zeros(12))
    assert abs(sum(b[0] for b in bodies) - 2.50000279) < 1e-8  # src/MPC.cpp:17
    assert np.allclose([C["BASE_IXX"], C["BASE_IYY"], C["BASE_IZZ"]], [0.00578574, 0.01938108, 0.02476124])  # MPC_crocoddyl.py:32
    assert np.allclose(np.abs(feet[:, 0]), 0.1946) and np.allclose(np.abs(feet[:, 1]), 0.14695)  # Controller.py:131-133, InvKin.cpp:9
    assert np.allclose(np.sign(feet[:, 0]), [1, 1, -1, -1]) and np.allclose(np.sign(feet[:, 1]), [1, -1, 1, -1])
    _, feet = _links(C, Q_INIT)
    assert np.allclose(feet[:, 2], -0.24474949993103629, atol=1e-12)  # scripts/test_mpc.py:39-40
    assert np.allclose(np.abs(feet[:, 0]), 0.1946, atol=1e-12)


def test_composite_inertia_and_centre_of_mass_reproduce_the_mpc_constants():
    """src/MPC.cpp:25-26 (gI) and :21 (offset_CoM) were derived from the real URDF; the header reproduces them."""
    C = _model()
    m, c, I = _composite(_links(C, Q_INIT)[0])
    rel = np.abs(np.diag(I) - np.diag(GI_REF)) / np.diag(GI_REF)
    assert rel.max() < 0.04, f"composite inertia {np.diag(I)} vs gI {np.diag(GI_REF)}: {rel}"
    off = I - np.diag(np.diag(I))
    assert np.abs(off).max() < 2e-4  # the reference's off-diagonal terms are 1e-4 and below
    assert abs(c[0]) < 1e-3 and abs(c[1]) < 1e-3
    assert -0.035 < c[2] < -0.02, f"centre of mass {c[2]:.4f} m under the base; the MPC assumes -0.03 (src/MPC.cpp:21)"


def test_oracle_crba_equals_the_numpy_rebuild(oracle):
    """Second implementation of the composite rigid body: the oracle's crba(q)[:6, :6] (C++) against the lines above."""
    C = _model()
    rng = np.random.default_rng(5)
    for qj in (np.zeros(12), Q_INIT, rng.uniform(-1.0, 1.0, 12)):
        q = np.zeros(19)
        q[6], q[7:] = 1.0, qj
        M = oracle.solo12_crba(q)
        m, c, I = _composite(_links(C, qj)[0])
        assert abs(M[0, 0] - m) < 1e-12
        Ib = I + m * (c @ c * np.eye(3) - np.outer(c, c))  # about the base origin
        np.testing.assert_allclose(M[3:6, 3:6], Ib, rtol=0, atol=1e-12)
        # first moment: the coupling block is +-m skew(c)
        S = M[0:3, 3:6]
        got = np.abs(np.array([S[2, 1], S[0, 2], S[1, 0]])) / m
        np.testing.assert_allclose(got, np.abs(c), atol=1e-12)
