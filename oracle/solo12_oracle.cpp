// oracle/solo12_oracle.cpp
//
// TEST INFRASTRUCTURE ONLY -- CPU oracle for the rigid-body quantities the reference obtains from Pinocchio
// (/root/reference/scripts/solo12InvKin.py:47-59, scripts/QP_WBC.py:89-117): foot placements / velocities /
// accelerations / Jacobians of the fixed-base model, contact Jacobians, crba diagonal and rnea of the
// floating-base model.  Pinocchio and the Solo12 URDF are absent from this container, so this is a
// restatement of the textbook algorithms Pinocchio implements (Featherstone, "Rigid Body Dynamics Algorithms":
// spatial-vector RNEA, table 5.1) over a generic parent-array tree with the constants of include/solo12_model.h.
// PARITY UNPINNED: no Pinocchio output exists here to compare with; tests pin it through identities instead
// (Jacobian = finite difference of the placement, rnea linear in the acceleration with a symmetric positive
// definite matrix whose base block carries the total mass, gravity wrench = weight at rest).
// It deliberately shares no code with the product (csrc/solo12_core.cuh works leg by leg with 3-vectors in the
// base frame; this file uses 6-D spatial vectors in link coordinates).
#include <cmath>
#include <cstring>

#include "../include/solo12_model.h"

namespace {

struct Vec3 {
  double v[3];
  double& operator[](int i) { return v[i]; }
  double operator[](int i) const { return v[i]; }
};
struct Mat3 {
  double m[3][3];
};
Vec3 mk(double x, double y, double z) { return Vec3{{x, y, z}}; }
Vec3 add(Vec3 a, Vec3 b) { return mk(a[0] + b[0], a[1] + b[1], a[2] + b[2]); }
Vec3 sub(Vec3 a, Vec3 b) { return mk(a[0] - b[0], a[1] - b[1], a[2] - b[2]); }
Vec3 scl(double s, Vec3 a) { return mk(s * a[0], s * a[1], s * a[2]); }
Vec3 crs(Vec3 a, Vec3 b) { return mk(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]); }
Vec3 mv(const Mat3& A, Vec3 x) {
  Vec3 r;
  for (int i = 0; i < 3; ++i) r[i] = A.m[i][0] * x[0] + A.m[i][1] * x[1] + A.m[i][2] * x[2];
  return r;
}
Vec3 mtv(const Mat3& A, Vec3 x) {
  Vec3 r;
  for (int i = 0; i < 3; ++i) r[i] = A.m[0][i] * x[0] + A.m[1][i] * x[1] + A.m[2][i] * x[2];
  return r;
}
Mat3 mm(const Mat3& A, const Mat3& B) {
  Mat3 C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[i][0] * B.m[0][j] + A.m[i][1] * B.m[1][j] + A.m[i][2] * B.m[2][j];
  return C;
}
Mat3 ident() { return Mat3{{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}}; }

// spatial motion / force vectors, Featherstone order: angular part first
struct Sv {
  Vec3 w, v;
};
// Pluecker transform parent -> child: E = rotation taking parent coordinates to child coordinates, r = position of the
// child origin in parent coordinates
struct Xf {
  Mat3 E;
  Vec3 r;
};
Sv xmot(const Xf& X, const Sv& m) { return Sv{mv(X.E, m.w), mv(X.E, sub(m.v, crs(X.r, m.w)))}; }
// force from child coordinates back to the parent: X' f
Sv xforceT(const Xf& X, const Sv& f) {
  const Vec3 fl = mtv(X.E, f.v);
  return Sv{add(mtv(X.E, f.w), crs(X.r, fl)), fl};
}
Sv crm(const Sv& a, const Sv& b) { return Sv{crs(a.w, b.w), add(crs(a.w, b.v), crs(a.v, b.w))}; }
Sv crf(const Sv& a, const Sv& f) { return Sv{add(crs(a.w, f.w), crs(a.v, f.v)), crs(a.w, f.v)}; }

struct Body {
  int parent;      // -1 for the base
  int axis;        // 0 = x, 1 = y (revolute); unused for the base
  Vec3 origin;     // joint placement in the parent frame
  double mass;
  Vec3 com;
  double I[6];     // xx xy xz yy yz zz about the centre of mass, link frame
};
Sv inertia_mul(const Body& b, const Sv& m) {
  // I_spatial [w; v] about the link origin: f = m (v + w x c)... standard form with h = m c
  const Vec3 c = b.com;
  const Vec3 vc = add(m.v, crs(m.w, c));  // velocity of the centre of mass
  const Vec3 lin = scl(b.mass, vc);
  const Vec3 Iw = mk(b.I[0] * m.w[0] + b.I[1] * m.w[1] + b.I[2] * m.w[2], b.I[1] * m.w[0] + b.I[3] * m.w[1] + b.I[4] * m.w[2],
                     b.I[2] * m.w[0] + b.I[4] * m.w[1] + b.I[5] * m.w[2]);
  return Sv{add(Iw, crs(c, lin)), lin};
}

struct Model {
  Body b[13];
  Vec3 foot[4];  // foot point in the lower-leg frame
  Model() {
    b[0] = Body{-1, 0, mk(0, 0, 0), S12_BASE_MASS, mk(0, 0, 0), {S12_BASE_IXX, 0, 0, S12_BASE_IYY, 0, S12_BASE_IZZ}};
    for (int leg = 0; leg < 4; ++leg) {
      const double sx = leg < 2 ? 1.0 : -1.0, sy = (leg & 1) ? -1.0 : 1.0;
      const int i0 = 1 + 3 * leg;
      b[i0] = Body{0, 0, mk(sx * S12_HAA_X, sy * S12_HAA_Y, 0), S12_L0_MASS, mk(sx * S12_L0_CX, sy * S12_L0_CY, S12_L0_CZ),
                   {S12_L0_IXX, sx * sy * S12_L0_IXY, sx * S12_L0_IXZ, S12_L0_IYY, sy * S12_L0_IYZ, S12_L0_IZZ}};
      b[i0 + 1] = Body{i0, 1, mk(0, sy * S12_HFE_Y, 0), S12_L1_MASS, mk(S12_L1_CX, sy * S12_L1_CY, S12_L1_CZ),
                       {S12_L1_IXX, sy * S12_L1_IXY, S12_L1_IXZ, S12_L1_IYY, sy * S12_L1_IYZ, S12_L1_IZZ}};
      b[i0 + 2] = Body{i0 + 1, 1, mk(0, sy * S12_KFE_Y, S12_KFE_Z), S12_L2_MASS, mk(S12_L2_CX, sy * S12_L2_CY, S12_L2_CZ),
                       {S12_L2_IXX, sy * S12_L2_IXY, S12_L2_IXZ, S12_L2_IYY, sy * S12_L2_IYZ, S12_L2_IZZ}};
      foot[leg] = mk(0, sy * S12_FOOT_Y, S12_FOOT_Z);
    }
  }
};
const Model& model() {
  static Model m;
  return m;
}
Xf joint_xf(const Body& b, double q) {
  const double c = std::cos(q), s = std::sin(q);
  Xf X;
  X.r = b.origin;
  if (b.axis == 0)
    X.E = Mat3{{{1, 0, 0}, {0, c, s}, {0, -s, c}}};
  else
    X.E = Mat3{{{c, 0, -s}, {0, 1, 0}, {s, 0, c}}};
  return X;
}
Mat3 quat_rot(const double* q) {  // (x, y, z, w) -> R (base to world)
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  return Mat3{{{1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)},
               {2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)},
               {2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)}}};
}

// rnea of the floating-base model; q[19], v[18], a[18] -> tau[18] in Pinocchio's layout (base: force then moment)
void rnea(const double* q, const double* v, const double* a, double* tau) {
  const Model& M = model();
  const Mat3 R = quat_rot(q + 3);
  Sv vel[13], acc[13], frc[13];
  Xf X[13];
  vel[0] = Sv{mk(v[3], v[4], v[5]), mk(v[0], v[1], v[2])};
  const Vec3 g = mtv(R, mk(0, 0, -S12_GRAVITY));
  acc[0] = Sv{mk(a[3], a[4], a[5]), sub(mk(a[0], a[1], a[2]), g)};
  for (int i = 1; i < 13; ++i) {
    const Body& b = M.b[i];
    X[i] = joint_xf(b, q[7 + i - 1]);
    Sv S{mk(b.axis == 0, b.axis == 1, 0), mk(0, 0, 0)};
    const Sv vj{scl(v[6 + i - 1], S.w), mk(0, 0, 0)};
    vel[i] = xmot(X[i], vel[b.parent]);
    vel[i].w = add(vel[i].w, vj.w);
    acc[i] = xmot(X[i], acc[b.parent]);
    const Sv cr = crm(vel[i], vj);
    acc[i].w = add(add(acc[i].w, scl(a[6 + i - 1], S.w)), cr.w);
    acc[i].v = add(acc[i].v, cr.v);
  }
  for (int i = 0; i < 13; ++i) {
    const Sv Ia = inertia_mul(M.b[i], acc[i]);
    const Sv Iv = inertia_mul(M.b[i], vel[i]);
    const Sv c = crf(vel[i], Iv);
    frc[i] = Sv{add(Ia.w, c.w), add(Ia.v, c.v)};
  }
  for (int i = 12; i >= 1; --i) {
    tau[6 + i - 1] = frc[i].w[M.b[i].axis];
    const Sv fp = xforceT(X[i], frc[i]);
    const int p = M.b[i].parent;
    frc[p].w = add(frc[p].w, fp.w);
    frc[p].v = add(frc[p].v, fp.v);
  }
  for (int c = 0; c < 3; ++c) {
    tau[c] = frc[0].v[c];
    tau[3 + c] = frc[0].w[c];
  }
}

// placement of the foot of `leg` in the base frame for joint angles qj[12]: position p and link orientation Rl (link -> base)
void foot_placement(const double* qj, int leg, Vec3& p, Mat3& Rl) {
  const Model& M = model();
  Mat3 Rb = ident();  // link -> base
  Vec3 o = mk(0, 0, 0);
  for (int l = 0; l < 3; ++l) {
    const Body& b = M.b[1 + 3 * leg + l];
    const Xf X = joint_xf(b, qj[3 * leg + l]);
    o = add(o, mv(Rb, b.origin));
    // X.E maps parent -> child, so child -> parent is its transpose
    Mat3 Et;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) Et.m[i][j] = X.E.m[j][i];
    Rb = mm(Rb, Et);
  }
  p = add(o, mv(Rb, M.foot[leg]));
  Rl = Rb;
}
// spatial velocity / acceleration (zero joint acceleration, fixed base, no gravity) of the lower leg of `leg`
void lower_leg_motion(const double* qj, const double* dqj, int leg, Sv& vel, Sv& acc) {
  const Model& M = model();
  Sv v{mk(0, 0, 0), mk(0, 0, 0)}, a{mk(0, 0, 0), mk(0, 0, 0)};
  for (int l = 0; l < 3; ++l) {
    const Body& b = M.b[1 + 3 * leg + l];
    const Xf X = joint_xf(b, qj[3 * leg + l]);
    const Sv vj{scl(dqj[3 * leg + l], mk(b.axis == 0, b.axis == 1, 0)), mk(0, 0, 0)};
    v = xmot(X, v);
    v.w = add(v.w, vj.w);
    a = xmot(X, a);
    const Sv cr = crm(v, vj);
    a.w = add(a.w, cr.w);
    a.v = add(a.v, cr.v);
  }
  vel = v;
  acc = a;
}

}  // namespace

extern "C" {

// scripts/solo12InvKin.py:47-59 for B robots; outputs in the layout the reference hands to InvKin (4 x 3 and 12 x 12
// column-major)
void orc_solo12_feet(int B, const double* q12, const double* dq12, double* posf, double* vf, double* wf, double* af,
                     double* Jf) {
  const Model& M = model();
  for (int b = 0; b < B; ++b) {
    const double* q = q12 + 12 * b;
    const double* dq = dq12 + 12 * b;
    double* J = Jf + 144 * b;
    std::memset(J, 0, 144 * sizeof(double));
    for (int leg = 0; leg < 4; ++leg) {
      Vec3 p;
      Mat3 Rl;
      foot_placement(q, leg, p, Rl);
      Sv v, a;
      lower_leg_motion(q, dq, leg, v, a);
      // frame at the foot, LOCAL_WORLD_ALIGNED: translate the spatial vectors to the foot point, rotate to base axes
      const Vec3 vl = mv(Rl, add(v.v, crs(v.w, M.foot[leg])));
      const Vec3 wl = mv(Rl, v.w);
      const Vec3 al = mv(Rl, add(a.v, crs(a.w, M.foot[leg])));
      for (int c = 0; c < 3; ++c) {
        posf[12 * b + leg + 4 * c] = p[c];
        vf[12 * b + leg + 4 * c] = vl[c];
        wf[12 * b + leg + 4 * c] = wl[c];
        af[12 * b + leg + 4 * c] = al[c];
      }
      // Jacobian column = foot velocity for a unit joint velocity
      for (int l = 0; l < 3; ++l) {
        double e[12] = {0};
        e[3 * leg + l] = 1.0;
        Sv v1, a1;
        lower_leg_motion(q, e, leg, v1, a1);
        const Vec3 col = mv(Rl, add(v1.v, crs(v1.w, M.foot[leg])));
        for (int c = 0; c < 3; ++c) J[(3 * leg + c) + 12 * (3 * leg + l)] = col[c];
      }
    }
  }
}
void orc_solo12_rnea(int B, const double* q, const double* v, const double* a, double* tau) {
  for (int b = 0; b < B; ++b) rnea(q + 19 * b, v + 18 * b, a + 18 * b, tau + 18 * b);
}
// scripts/QP_WBC.py:96-101: rows of the floating-base frame Jacobians (LOCAL_WORLD_ALIGNED, linear) of feet in contact
void orc_solo12_jc(int B, const double* q, const double* contacts, double* Jc) {
  for (int b = 0; b < B; ++b) {
    const double* qb = q + 19 * b;
    double* J = Jc + 216 * b;
    std::memset(J, 0, 216 * sizeof(double));
    const Mat3 R = quat_rot(qb + 3);
    double Jf[144], posf[12], t1[12], t2[12], t3[12], zero[12] = {0};
    orc_solo12_feet(1, qb + 7, zero, posf, t1, t2, t3, Jf);
    for (int leg = 0; leg < 4; ++leg) {
      if (contacts[4 * b + leg] == 0.0) continue;
      const Vec3 p = mk(posf[leg], posf[leg + 4], posf[leg + 8]);
      for (int c = 0; c < 18; ++c) {
        Vec3 col = mk(0, 0, 0);
        if (c < 3)
          col = mk(R.m[0][c], R.m[1][c], R.m[2][c]);  // base linear velocity is expressed in the base frame
        else if (c < 6) {
          Vec3 e = mk(c == 3, c == 4, c == 5);
          col = mv(R, crs(e, p));
        } else
          col = mv(R, mk(Jf[(3 * leg) + 12 * (c - 6)], Jf[(3 * leg + 1) + 12 * (c - 6)], Jf[(3 * leg + 2) + 12 * (c - 6)]));
        for (int r = 0; r < 3; ++r) J[(3 * leg + r) + 12 * c] = col[r];
      }
    }
  }
}
// crba by columns of rnea at zero velocity without gravity: M e_j = rnea(q, 0, e_j) - rnea(q, 0, 0); 18 x 18 column-major
void orc_solo12_crba(const double* q, double* Mout) {
  double v[18] = {0}, a[18] = {0}, t0[18], t1[18];
  rnea(q, v, a, t0);
  for (int j = 0; j < 18; ++j) {
    a[j] = 1.0;
    rnea(q, v, a, t1);
    a[j] = 0.0;
    for (int i = 0; i < 18; ++i) Mout[i + 18 * j] = t1[i] - t0[i];
  }
}

}  // extern "C"
