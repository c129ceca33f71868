// solo12_core.cuh -- Solo12 leg kinematics and recursive Newton-Euler dynamics, one LEG per thread.
//
// Replaces, for a batch of robots, the Pinocchio calls the reference makes from Python around its C++ classes
// (SURVEY.md section 8 row a18):
//   scripts/solo12InvKin.py:47-59   computeJointJacobians / forwardKinematics(q, dq, 0) / updateFramePlacements on the
//                                   fixed-base model; oMf.translation, getFrameVelocity / getFrameAcceleration /
//                                   getFrameJacobian (LOCAL_WORLD_ALIGNED) of the four feet
//   scripts/QP_WBC.py:89-93         crba at the zero configuration, 6x6 base block masked to its diagonal
//   scripts/QP_WBC.py:96-101        contact Jacobians of the floating-base model (LOCAL_WORLD_ALIGNED, linear rows)
//   scripts/QP_WBC.py:104,116-117   rnea(q, dq, ddq) base wrench and joint torques, tau_ff = rnea[6:] - Jc[:,6:]' f
// The robot is a tree of four identical 3-joint chains on a floating base, so a thread carries one chain with plain
// 3-vectors expressed in the base frame (classical Newton-Euler); the four threads of a robot combine their chains'
// reactions on the base by shuffles.  Model constants: include/solo12_model.h (PARITY UNPINNED, see there).
#pragma once
#include "../../include/solo12_model.h"
#include "qrw_spmd.cuh"

namespace qrw {
namespace s12 {

struct V3 {
  double x, y, z;
};
QRW_HD V3 v3(double x, double y, double z) { return V3{x, y, z}; }
QRW_HD V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
QRW_HD V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
QRW_HD V3 operator*(double s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
QRW_HD V3 cross(V3 a, V3 b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
QRW_HD double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct M3 {  // columns
  V3 c0, c1, c2;
};
QRW_HD V3 mul(const M3& R, V3 v) { return v.x * R.c0 + (v.y * R.c1 + v.z * R.c2); }
QRW_HD V3 mulT(const M3& R, V3 v) { return V3{dot(R.c0, v), dot(R.c1, v), dot(R.c2, v)}; }

// rotation of the base from the quaternion (x, y, z, w) Pinocchio keeps in q[3:7]
QRW_HD M3 quat_to_rot(double x, double y, double z, double w) {
  M3 R;
  R.c0 = v3(1.0 - 2.0 * (y * y + z * z), 2.0 * (x * y + z * w), 2.0 * (x * z - y * w));
  R.c1 = v3(2.0 * (x * y - z * w), 1.0 - 2.0 * (x * x + z * z), 2.0 * (y * z + x * w));
  R.c2 = v3(2.0 * (x * z + y * w), 2.0 * (y * z - x * w), 1.0 - 2.0 * (x * x + y * y));
  return R;
}

// symmetric inertia (link frame) rotated into the base frame and applied to a vector: R I R' v
struct Sym {
  double xx, xy, xz, yy, yz, zz;
};
QRW_HD V3 inertia_apply(const M3& R, const Sym& I, V3 v) {
  const V3 l = mulT(R, v);
  const V3 m = v3(I.xx * l.x + I.xy * l.y + I.xz * l.z, I.xy * l.x + I.yy * l.y + I.yz * l.z,
                  I.xz * l.x + I.yz * l.y + I.zz * l.z);
  return mul(R, m);
}

// Kinematics of one leg in the base frame: joint origins, joint axes, link orientations, foot point.
struct LegKin {
  M3 R0, R1, R2;  // orientation of hip assembly, upper leg, lower leg
  V3 o0, o1, o2;  // joint origins (HAA, HFE, KFE)
  V3 s1;          // axis of HFE and KFE (the HAA axis is x)
  V3 pf;          // foot point
};
// leg = 0..3 (FL, FR, HL, HR); q = the leg's three joint angles
QRW_HD LegKin leg_kinematics(int leg, const double* q) {
  const double sx = (leg < 2) ? 1.0 : -1.0, sy = (leg & 1) ? -1.0 : 1.0;
  LegKin K;
  const double c0 = cos(q[0]), s0 = sin(q[0]);
  const double c1 = cos(q[1]), s1 = sin(q[1]);
  const double c12 = cos(q[1] + q[2]), s12 = sin(q[1] + q[2]);
  K.R0.c0 = v3(1.0, 0.0, 0.0);
  K.R0.c1 = v3(0.0, c0, s0);
  K.R0.c2 = v3(0.0, -s0, c0);
  // R1 = R0 Ry(q1), R2 = R0 Ry(q1 + q2);  Ry(a) columns: (c, 0, -s), (0, 1, 0), (s, 0, c)
  K.R1.c0 = c1 * K.R0.c0 - s1 * K.R0.c2;
  K.R1.c1 = K.R0.c1;
  K.R1.c2 = s1 * K.R0.c0 + c1 * K.R0.c2;
  K.R2.c0 = c12 * K.R0.c0 - s12 * K.R0.c2;
  K.R2.c1 = K.R0.c1;
  K.R2.c2 = s12 * K.R0.c0 + c12 * K.R0.c2;
  K.s1 = K.R0.c1;
  K.o0 = v3(sx * S12_HAA_X, sy * S12_HAA_Y, 0.0);
  K.o1 = K.o0 + mul(K.R0, v3(0.0, sy * S12_HFE_Y, 0.0));
  K.o2 = K.o1 + mul(K.R1, v3(0.0, sy * S12_KFE_Y, S12_KFE_Z));
  K.pf = K.o2 + mul(K.R2, v3(0.0, sy * S12_FOOT_Y, S12_FOOT_Z));
  return K;
}
// columns of the 3x3 leg Jacobian d pf / d q (base frame)
QRW_HD void leg_jacobian(const LegKin& K, V3 J[3]) {
  J[0] = cross(v3(1.0, 0.0, 0.0), K.pf - K.o0);
  J[1] = cross(K.s1, K.pf - K.o1);
  J[2] = cross(K.s1, K.pf - K.o2);
}

// Velocity-level quantities of the foot for a FIXED base (scripts/solo12InvKin.py:47-59): linear velocity vf,
// angular velocity wf and the classical acceleration of the foot point with zero joint accelerations.
// (Pinocchio's getFrameAcceleration returns the spatial one, a_cl - wf x vf; InvKin.cpp:48 adds wf x vf back.)
QRW_HD void foot_motion(const LegKin& K, const double* dq, V3& vf, V3& wf, V3& acl) {
  const V3 ex = v3(1.0, 0.0, 0.0);
  const V3 w0 = dq[0] * ex;
  const V3 w1 = w0 + dq[1] * K.s1;
  const V3 w2 = w1 + dq[2] * K.s1;
  const V3 al1 = cross(w0, dq[1] * K.s1);             // alpha0 = 0 (base fixed, ddq = 0)
  const V3 al2 = al1 + cross(w1, dq[2] * K.s1);
  const V3 d01 = K.o1 - K.o0, d12 = K.o2 - K.o1, d2f = K.pf - K.o2;
  const V3 v1 = cross(w0, d01);
  const V3 v2 = v1 + cross(w1, d12);
  vf = v2 + cross(w2, d2f);
  wf = w2;
  const V3 a1 = cross(w0, cross(w0, d01));
  const V3 a2 = a1 + cross(al1, d12) + cross(w1, cross(w1, d12));
  acl = a2 + cross(al2, d2f) + cross(w2, cross(w2, d2f));
}

struct LegDyn {
  double tau[3];  // joint torques
  V3 f, n;        // force and moment (about the BASE ORIGIN) the chain needs from the base, base frame
};
// Recursive Newton-Euler of one chain.  wb, alb: angular velocity / acceleration of the base; ab: classical linear
// acceleration of the base origin minus gravity; all in the base frame.
QRW_HD LegDyn leg_dynamics(int leg, const LegKin& K, V3 wb, V3 alb, V3 ab, const double* dq, const double* ddq) {
  const double sx = (leg < 2) ? 1.0 : -1.0, sy = (leg & 1) ? -1.0 : 1.0;
  const V3 ex = v3(1.0, 0.0, 0.0);
  // link constants (mirrored)
  const V3 c0 = mul(K.R0, v3(sx * S12_L0_CX, sy * S12_L0_CY, S12_L0_CZ));
  const V3 c1 = mul(K.R1, v3(S12_L1_CX, sy * S12_L1_CY, S12_L1_CZ));
  const V3 c2 = mul(K.R2, v3(S12_L2_CX, sy * S12_L2_CY, S12_L2_CZ));
  const Sym I0{S12_L0_IXX, sx * sy * S12_L0_IXY, sx * S12_L0_IXZ, S12_L0_IYY, sy * S12_L0_IYZ, S12_L0_IZZ};
  const Sym I1{S12_L1_IXX, sy * S12_L1_IXY, S12_L1_IXZ, S12_L1_IYY, sy * S12_L1_IYZ, S12_L1_IZZ};
  const Sym I2{S12_L2_IXX, sy * S12_L2_IXY, S12_L2_IXZ, S12_L2_IYY, sy * S12_L2_IYZ, S12_L2_IZZ};
  // forward: angular velocities / accelerations, accelerations of the joint origins
  const V3 w0 = wb + dq[0] * ex;
  const V3 al0 = alb + ddq[0] * ex + cross(wb, dq[0] * ex);
  const V3 a0 = ab + cross(alb, K.o0) + cross(wb, cross(wb, K.o0));
  const V3 w1 = w0 + dq[1] * K.s1;
  const V3 al1 = al0 + ddq[1] * K.s1 + cross(w0, dq[1] * K.s1);
  const V3 d01 = K.o1 - K.o0;
  const V3 a1 = a0 + cross(al0, d01) + cross(w0, cross(w0, d01));
  const V3 w2 = w1 + dq[2] * K.s1;
  const V3 al2 = al1 + ddq[2] * K.s1 + cross(w1, dq[2] * K.s1);
  const V3 d12 = K.o2 - K.o1;
  const V3 a2 = a1 + cross(al1, d12) + cross(w1, cross(w1, d12));
  // inertial force / moment about the centre of mass of every link
  const V3 F0 = S12_L0_MASS * (a0 + cross(al0, c0) + cross(w0, cross(w0, c0)));
  const V3 F1 = S12_L1_MASS * (a1 + cross(al1, c1) + cross(w1, cross(w1, c1)));
  const V3 F2 = S12_L2_MASS * (a2 + cross(al2, c2) + cross(w2, cross(w2, c2)));
  const V3 N0 = inertia_apply(K.R0, I0, al0) + cross(w0, inertia_apply(K.R0, I0, w0));
  const V3 N1 = inertia_apply(K.R1, I1, al1) + cross(w1, inertia_apply(K.R1, I1, w1));
  const V3 N2 = inertia_apply(K.R2, I2, al2) + cross(w2, inertia_apply(K.R2, I2, w2));
  // backward: moments about the joint origins
  LegDyn D;
  const V3 f2 = F2;
  const V3 n2 = N2 + cross(c2, F2);
  D.tau[2] = dot(K.s1, n2);
  const V3 f1 = F1 + f2;
  const V3 n1 = N1 + cross(c1, F1) + n2 + cross(d12, f2);
  D.tau[1] = dot(K.s1, n1);
  const V3 f0 = F0 + f1;
  const V3 n0 = N0 + cross(c0, F0) + n1 + cross(d01, f1);
  D.tau[0] = n0.x;
  D.f = f0;
  D.n = n0 + cross(K.o0, f0);
  return D;
}
// the base link itself (centre of mass at the origin of the base frame)
QRW_HD void base_dynamics(V3 wb, V3 alb, V3 ab, V3& f, V3& n) {
  f = S12_BASE_MASS * ab;
  const V3 Iw = v3(S12_BASE_IXX * wb.x, S12_BASE_IYY * wb.y, S12_BASE_IZZ * wb.z);
  n = v3(S12_BASE_IXX * alb.x, S12_BASE_IYY * alb.y, S12_BASE_IZZ * alb.z) + cross(wb, Iw);
}

// 3x3 inverse by the adjugate (columns in, columns out): the InvKin inverse of a leg's Jacobian block
QRW_HD void inv3(const V3 J[3], V3 Ji[3]) {
  const V3 r0 = cross(J[1], J[2]), r1 = cross(J[2], J[0]), r2 = cross(J[0], J[1]);  // rows of adj
  const double id = 1.0 / dot(J[0], r0);
  Ji[0] = v3(r0.x * id, r1.x * id, r2.x * id);
  Ji[1] = v3(r0.y * id, r1.y * id, r2.y * id);
  Ji[2] = v3(r0.z * id, r1.z * id, r2.z * id);
}

}  // namespace s12
}  // namespace qrw
