// wbc_tick_core.cuh -- the whole-body tick (Solo12 kinematics / RNEA + InvKin + QPWBC + tau_ff) for EIGHT robots per warp.
//
// What it replaces (reference, CPU, one robot): wbc_controller.compute, /root/reference/scripts/QP_WBC.py:52-131, i.e.
//   Solo12InvKin.refreshAndCompute   scripts/solo12InvKin.py:47-59 + InvKin::refreshAndCompute src/InvKin.cpp:23-73
//   Pinocchio crba / contact Jacobians / rnea        scripts/QP_WBC.py:89-104
//   QPWBC::run                                        src/QPWBC.cpp:310-343 (compute_matrices :481-498, update_PQ :520-537,
//                                                     call_solver :213-275 -> OSQP, retrieve_result :277-297)
//   tau_ff = rnea(q, dq, ddq)[6:] - Jc[:, 6:]' f      scripts/QP_WBC.py:116-117
//
// Why a second layout (round 2).  ncu on the warp-per-robot kernel (profiles/r02p_ncu_wbc_tick.txt): instruction-bound
// (19.4 k warp instructions per robot, 11.5 of 32 lanes active on average): the leg code runs on 4 lanes, the 12-variable /
// 20-row QP on 12..20 lanes, and every cross-lane step is a shared-memory round trip with a warp barrier.  Here
//   * the leg code runs on all 32 lanes: lane = 4 * robot + leg for the eight robots of an OCTET;
//   * the QP runs on HALF-WARPS, two robots at a time (four rounds per octet).  Lane j < 12 of a half owns variable j
//     (foot f = j / 3, axis c = j % 3) AND the friction-cone rows of that axis: slot A = row 5f + 2c, slot B = row 5f + 2c + 1
//     (axis z: row 5f + 4 in slot A, no slot B).  x, z, y, D, E, bounds and q live in registers of their owner; G and G'
//     products are shuffles inside the foot's three lanes; the 12 x 12 inverse is a Gauss-Jordan on register rows with
//     the pivot row shuffled; only the 12-vectors every lane needs (right-hand side, D, x) go through shared memory;
//   * what the solver needs of the contact Jacobian is its six base columns (X = Jc[:, :6]'): 72 + 6 doubles per robot are
//     handed over on chip, nothing else.
// Same arithmetic, in the same order, as wbc_core.cuh (which stays for the standalone QPWBC entry points): same ADMM
// iteration counts, results equal to round-off.  Written against the per-thread layer of qrw_spmd.cuh, so the whole tick
// also runs in the CPU fiber emulation (tests/emu) against the oracle.
#pragma once
#include "solo12_core.cuh"
#include "wbc_core.cuh"

namespace qrw {
namespace wt {

constexpr int LDH = 14;  // padded row length: conflict-free LDS.128 of 12 different rows
// shared memory of a half-warp (doubles)
constexpr int P_HC = 0;     // [12][14] column j of P (the mirrored upper triangle of H, QPWBC.cpp:520-532)
constexpr int P_KI = 168;   // [12][14] row j of K^-1; until the first factorisation: [6][12] A * q1
constexpr int P_A = 336;    // [6][12]  A = Yinv X, row-major
constexpr int P_V0 = 408;   // [16] broadcast vector: f_cmd, second column norms, right-hand side, x, delta f
constexpr int P_V1 = 424;   // [16] broadcast vector: D
constexpr int P_V2 = 440;   // [16] broadcast vector: D inside the Ruiz passes
constexpr int P_GAM = 456;  // [8]  gamma
constexpr int P_XF = 464;   // [8]  X f_cmd - RNEA
constexpr int P_DOUBLES = 472;
// shared memory of a warp
constexpr int T_JC6 = 0;     // [8][72] base columns of the contact Jacobian of the octet's robots (12 x 6, column-major)
constexpr int T_RNEA = 576;  // [8][8]  base wrench of rnea(q, dq, [0; ddq_cmd])
constexpr int T_HALF = 640;  // two solver slabs
constexpr int WT2_SLAB_DOUBLES = T_HALF + 2 * P_DOUBLES;
constexpr int WT2_SLAB_BYTES = WT2_SLAB_DOUBLES * 8;
static_assert(P_KI % 2 == 0 && P_A % 2 == 0 && P_V0 % 2 == 0 && P_DOUBLES % 2 == 0 && T_HALF % 2 == 0, "16-byte alignment");

// max over the 16 lanes of a half-warp, on every lane (OSQP's vec_norm_inf: a NaN never wins)
QRW_HD double hmax16(double v) {
#pragma unroll
  for (int m = 8; m > 0; m >>= 1) {
    const double o = cta::hxor(v, m);
    v = o > v ? o : v;
  }
  return v;
}
QRW_HD double amax(double m, double v) {
  const double a = fabs(v);
  return a > m ? a : m;
}
// (G v)[row] for the rows of a lane: v = the lane's entry, vz = the entry of its foot's z lane (QPWBC.cpp:4-22)
QRW_HD double g_rowA(double mu, int c, double v, double vz) { return c == 2 ? v : (-v) + mu * vz; }
QRW_HD double g_rowB(double mu, double v, double vz) { return v + mu * vz; }
// (G' w)[j]: wA, wB = the lane's two row entries
QRW_HD double gt_col(double mu, int f, int c, double wA, double wB) {
  const double a0 = cta::hshfl(wA, 3 * f), b0 = cta::hshfl(wB, 3 * f);
  const double a1 = cta::hshfl(wA, 3 * f + 1), b1 = cta::hshfl(wB, 3 * f + 1);
  return c == 2 ? mu * (a0 + b0 + a1 + b1) + wA : wB - wA;
}
// entry of G in row 5f + rw, column 3f + ci
QRW_HD double g_ent(double mu, int rw, int ci) {
  if (rw == 4) return ci == 2 ? 1.0 : 0.0;
  if (ci == 2) return mu;
  if (ci == (rw >> 1)) return (rw & 1) ? 1.0 : -1.0;
  return 0.0;
}

// registers of a lane across the solve
struct Lane {
  double x, D, qh, dinv;
  double zA, yA, eA, loA, upA, einvA, rrA, riA;
  double zB, yB, eB, loB, upB, einvB, rrB, riB;
};

// K = P_hat + sigma I + A_hat' diag(rho) A_hat (row j in registers), inverted by Gauss-Jordan with the pivot row shuffled;
// row j of K^-1 -> shared memory.  Same operations as wbc_factor (wbc_core.cuh).  V1 holds D.
QRW_HDN void factor_pair(const double sigma, const double mu, double* ps, const double D, const double eA, const double eB, const double rrA,
                         const double rrB, const double cscale, const int j, const int f, const int c, const bool isv) {
  double kr[12], Dv[12], hc[12];
  cta::lds12(ps + P_V1, Dv);
  cta::lds12(ps + P_HC + LDH * j, hc);
#pragma unroll
  for (int cc = 0; cc < 12; ++cc) {
    double v = cscale * D * hc[cc] * Dv[cc];
    if (cc == j) v += sigma;
    kr[cc] = v;
  }
  // the foot's 3 x 3 block of A' diag(rho) A: rows 5f .. 5f + 4 sit in slots A, B of lanes 3f, 3f + 1 and slot A of lane 3f + 2
  double e5[5], r5[5];
  e5[0] = cta::hshfl(eA, 3 * f), r5[0] = cta::hshfl(rrA, 3 * f);
  e5[1] = cta::hshfl(eB, 3 * f), r5[1] = cta::hshfl(rrB, 3 * f);
  e5[2] = cta::hshfl(eA, 3 * f + 1), r5[2] = cta::hshfl(rrA, 3 * f + 1);
  e5[3] = cta::hshfl(eB, 3 * f + 1), r5[3] = cta::hshfl(rrB, 3 * f + 1);
  e5[4] = cta::hshfl(eA, 3 * f + 2), r5[4] = cta::hshfl(rrA, 3 * f + 2);
  double acc3[3];
#pragma unroll
  for (int c2 = 0; c2 < 3; ++c2) {
    double acc = 0.0;
#pragma unroll
    for (int rw = 0; rw < 5; ++rw) acc += r5[rw] * (e5[rw] * g_ent(mu, rw, c)) * (e5[rw] * g_ent(mu, rw, c2));
    acc3[c2] = acc;
  }
#pragma unroll
  for (int cc = 0; cc < 12; ++cc)
    if (cc / 3 == f) kr[cc] += acc3[cc % 3] * D * Dv[cc];
#pragma unroll
  for (int p = 0; p < 12; ++p) {
    double pr[12];
#pragma unroll
    for (int cc = 0; cc < 12; ++cc) pr[cc] = cta::hshfl(kr[cc], p);
    const double d = 1.0 / pr[p];
    const double colp = kr[p];
#pragma unroll
    for (int cc = 0; cc < 12; ++cc) {
      const double t = (cc == p) ? d : pr[cc] * d;
      double v;
      if (j == p)
        v = t;
      else if (cc == p)
        v = -colp * d;
      else
        v = kr[cc] - colp * t;
      kr[cc] = v;
    }
  }
  if (isv) {
#pragma unroll
    for (int cc = 0; cc < 12; ++cc) ps[P_KI + LDH * j + cc] = kr[cc];
  }
  cta::wsync();
}

// One QPWBC::run for the robot of this half-warp.  hl = lane & 15.  `valid` = false: the half has no robot (the tail of the
// batch): it runs the same instruction stream on zeros and stores nothing.
// Jc6 [12 x 6] column-major and rnea [6]: shared memory; f_cmd, st, Yinv, f_res, ddq_res, info: global memory of the robot.
QRW_HD void solve_pair(const WbcConst& C, double* ps, const int hl, const bool valid, const double* Jc6, const double* rnea,
                        const double* f_cmd, double* st, const int first, const double* __restrict__ Yinv, double* f_res,
                        double* ddq_res, int* info) {
  const bool isv = hl < 12;
  const int j = isv ? hl : 11;  // lanes 12..15 shadow lane 11 and store nothing
  const int f = j / 3, c = j - 3 * f;
  const bool hasB = c < 2;
  const int rA = 5 * f + 2 * c, rB = rA + 1;
  const bool warm = valid && !first;
  Lane L;
  L.x = warm ? st[WS_X + j] : 0.0;
  L.zA = warm ? st[WS_Z + rA] : 0.0;
  L.yA = warm ? st[WS_Y + rA] : 0.0;
  L.zB = (warm && hasB) ? st[WS_Z + rB] : 0.0;
  L.yB = (warm && hasB) ? st[WS_Y + rB] : 0.0;
  const double fc = valid ? f_cmd[j] : 0.0;
  // ---- compute_matrices (QPWBC.cpp:481-498): A = Yinv X with X[r][c] = Jc(c, r); lane j owns column j of A
  double acol[6], aq[6];
#pragma unroll
  for (int r = 0; r < 6; ++r) {
    double s = 0;
#pragma unroll
    for (int k = 0; k < 6; ++k) s += Yinv[r + 6 * k] * Jc6[j + 12 * k];
    acol[r] = s;
    aq[r] = s * C.q1;
  }
  double* AQ = ps + P_KI;
  if (isv) {
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      ps[P_A + 12 * r + j] = acol[r];
      AQ[12 * r + j] = aq[r];
    }
    ps[P_V0 + j] = fc;
  }
  cta::wsync();
  if (hl < 6) {
    double fv[12], s = 0;
    cta::lds12(ps + P_V0, fv);
#pragma unroll
    for (int cc = 0; cc < 12; ++cc) s += Jc6[cc + 12 * hl] * fv[cc];
    ps[P_XF + hl] = s - rnea[hl];
  }
  cta::wsync();
  if (hl < 6) {
    double s = 0;
#pragma unroll
    for (int k = 0; k < 6; ++k) s += Yinv[hl + 6 * k] * ps[P_XF + k];
    ps[P_GAM + hl] = s;
  }
  // column j of P: P(i, j) = H(i, j) for i <= j, H(j, i) otherwise, H(i, j) = sum_k (A(k, i) q1) A(k, j) + q2 [i == j]
  double hc[12];
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    double su[6], sl[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) su[i] = 0.0, sl[i] = 0.0;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      double a[6], q[6];
      cta::lds6(ps + P_A + 12 * k + 6 * g, a);
      cta::lds6(AQ + 12 * k + 6 * g, q);
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        su[i] += q[i] * acol[k];
        sl[i] += aq[k] * a[i];
      }
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const int ii = 6 * g + i;
      hc[ii] = (ii <= j ? su[i] : sl[i]) + (ii == j ? C.q2 : 0.0);
    }
  }
  cta::wsync();  // gamma is visible; everybody is done with A * q1 (the K^-1 slot)
  double g_ = 0;
#pragma unroll
  for (int k = 0; k < 6; ++k) g_ += aq[k] * ps[P_GAM + k];
  // osqp_update_P rescales with the q still held by the workspace (the previous call's g)
  const double qold = warm ? st[WS_Q + j] : g_;
  if (isv) {
#pragma unroll
    for (int i = 0; i < 12; ++i) ps[P_HC + LDH * j + i] = hc[i];
  }
  {  // bounds (QPWBC.cpp:337-343), unscaled for now
    const double fz = cta::hshfl(fc, 3 * f + 2);
    const double GfA = g_rowA(C.mu, c, fc, fz), GfB = g_rowB(C.mu, fc, fz);
    L.loA = -GfA, L.upA = -GfA + C.fmax;
    L.loB = -GfB, L.upB = -GfB + C.fmax;
  }
  // ---- Ruiz equilibration (OSQP scale_data) on P, A = G, q = qold
  double cscale = 1.0;
  L.D = 1.0, L.eA = 1.0, L.eB = 1.0;
#pragma unroll 1
  for (int pass = 0; pass < 10; ++pass) {
    double Dv[12];
    if (isv) ps[P_V1 + j] = L.D;
    cta::wsync();
    cta::lds12(ps + P_V1, Dv);
    // max_i |(cscale D_i h_i) D_j| = (max_i |cscale D_i h_i|) D_j to the bit: D_j > 0 and rounding is monotone
    double nrm = 0.0;
#pragma unroll
    for (int i = 0; i < 12; ++i) nrm = amax(nrm, cscale * Dv[i] * hc[i]);
    nrm *= L.D;
    const double eA0 = cta::hshfl(L.eA, 3 * f), eB0 = cta::hshfl(L.eB, 3 * f);
    const double eA1 = cta::hshfl(L.eA, 3 * f + 1), eB1 = cta::hshfl(L.eB, 3 * f + 1);
    const double Dz = cta::hshfl(L.D, 3 * f + 2);
    double nA, nB;
    if (c == 2) {
      nrm = amax(nrm, eA0 * C.mu * L.D);
      nrm = amax(nrm, eB0 * C.mu * L.D);
      nrm = amax(nrm, eA1 * C.mu * L.D);
      nrm = amax(nrm, eB1 * C.mu * L.D);
      nrm = amax(nrm, L.eA * L.D);
      nA = fabs(L.eA * L.D);
      nB = 0.0;
    } else {
      nrm = amax(nrm, L.eA * L.D);
      nrm = amax(nrm, L.eB * L.D);
      nA = amax(fabs(L.eA * L.D), L.eA * C.mu * Dz);
      nB = amax(fabs(L.eB * L.D), L.eB * C.mu * Dz);
    }
    const double dfac = 1.0 / sqrt(limit_scaling(nrm));
    const double efA = 1.0 / sqrt(limit_scaling(nA)), efB = 1.0 / sqrt(limit_scaling(nB));
    L.D *= dfac;
    L.eA *= efA;
    L.eB *= efB;
    if (isv) ps[P_V2 + j] = L.D;
    cta::wsync();
    cta::lds12(ps + P_V2, Dv);
    double n2 = 0.0;
#pragma unroll
    for (int i = 0; i < 12; ++i) n2 = amax(n2, cscale * Dv[i] * hc[i]);
    n2 *= L.D;
    if (isv) ps[P_V0 + j] = n2;
    double qn = hmax16(fabs(cscale * L.D * qold));
    cta::wsync();
    double nv[12], mean = 0.0;
    cta::lds12(ps + P_V0, nv);
#pragma unroll
    for (int jj = 0; jj < 12; ++jj) mean += nv[jj];
    mean /= 12.0;
    qn = limit_scaling(qn);
    const double c_temp = limit_scaling(dmax(mean, qn));
    cscale *= 1.0 / c_temp;
  }
  const double cinv = 1.0 / cscale;
  L.qh = cscale * (L.D * g_);  // osqp_update_lin_cost
  L.loA = L.eA * L.loA, L.upA = L.eA * L.upA;
  L.loB = L.eB * L.loB, L.upB = L.eB * L.upB;
  L.dinv = 1.0 / L.D, L.einvA = 1.0 / L.eA, L.einvB = 1.0 / L.eB;
  if (isv) ps[P_V1 + j] = L.D;  // stays there for the factorisations and the residual checks
  cta::wsync();
  double rho = warm ? st[WS_RHO] : C.rho0;
  rho = dmin(dmax(rho, 1e-6), 1e6);
  const bool eqA = (L.upA - L.loA) < 1e-4, eqB = (L.upB - L.loB) < 1e-4;
  L.rrA = eqA ? 1e3 * rho : rho, L.rrB = eqB ? 1e3 * rho : rho;
  L.riA = 1.0 / L.rrA, L.riB = 1.0 / L.rrB;
  int n_factor = 1;
  factor_pair(C.sigma, C.mu, ps, L.D, L.eA, L.eB, L.rrA, L.rrB, cscale, j, f, c, isv);

  // ---- ADMM.  The two halves of the warp iterate in lockstep; a half that has terminated keeps its iterates frozen.
  int status = -10, iter_done = 0;
  bool act = true;
  int until_check = C.check_every, until_adapt = C.adapt_every;  // countdowns instead of iter % n (20 instructions each)
#pragma unroll 1
  for (int iter = 1; iter <= C.max_iter; ++iter) {
    const double wA = L.eA * (L.rrA * L.zA - L.yA);
    const double wB = hasB ? L.eB * (L.rrB * L.zB - L.yB) : 0.0;
    const double gt = gt_col(C.mu, f, c, wA, wB);
    const double rhs = C.sigma * L.x - L.qh + L.D * gt;
    if (isv) ps[P_V0 + j] = rhs;
    cta::wsync();
    double rv[12], kr[12];
    cta::lds12(ps + P_V0, rv);
    cta::lds12(ps + P_KI + LDH * j, kr);
    double s0 = 0, s1 = 0;
#pragma unroll
    for (int i = 0; i < 12; i += 2) {
      s0 += kr[i] * rv[i];
      s1 += kr[i + 1] * rv[i + 1];
    }
    const double xt = s0 + s1;
    const double dx = L.D * xt;
    const double dz = cta::hshfl(dx, 3 * f + 2);  // also orders this iteration's reads of V0 before the next one's write
    {
      const double zt = L.eA * g_rowA(C.mu, c, dx, dz);
      const double zr = C.alpha * zt + (1.0 - C.alpha) * L.zA;
      const double zn = dmin(dmax(zr + L.riA * L.yA, L.loA), L.upA);
      if (act) {
        L.yA += L.rrA * (zr - zn);
        L.zA = zn;
      }
    }
    if (hasB) {
      const double zt = L.eB * g_rowB(C.mu, dx, dz);
      const double zr = C.alpha * zt + (1.0 - C.alpha) * L.zB;
      const double zn = dmin(dmax(zr + L.riB * L.yB, L.loB), L.upB);
      if (act) {
        L.yB += L.rrB * (zr - zn);
        L.zB = zn;
      }
    }
    if (act) L.x = C.alpha * xt + (1.0 - C.alpha) * L.x;
    --until_adapt;
    if (--until_check == 0) {
      until_check = C.check_every;
      // residuals (auxil.c compute_pri_res / compute_dua_res / compute_*_tol)
      cta::wsync();
      if (isv) ps[P_V0 + j] = L.x;
      cta::wsync();
      double xv[12], Dv[12], hc2[12];
      cta::lds12(ps + P_V0, xv);
      cta::lds12(ps + P_V1, Dv);
      cta::lds12(ps + P_HC + LDH * j, hc2);
      const double dxx = L.D * L.x;
      const double dzz = cta::hshfl(dxx, 3 * f + 2);
      const double axA = L.eA * g_rowA(C.mu, c, dxx, dzz);
      const double axB = hasB ? L.eB * g_rowB(C.mu, dxx, dzz) : 0.0;
      double px = 0.0;
#pragma unroll
      for (int i = 0; i < 12; ++i) px += (cscale * Dv[i] * hc2[i] * L.D) * xv[i];
      const double aty = L.D * gt_col(C.mu, f, c, L.eA * L.yA, hasB ? L.eB * L.yB : 0.0);
      const double rsA = axA - L.zA, rsB = axB - L.zB;
      double pri_u = amax(fabs(L.einvA * rsA), L.einvB * rsB);
      double relp = amax(amax(fabs(L.einvA * L.zA), L.einvB * L.zB), dmax(fabs(L.einvA * axA), fabs(L.einvB * axB)));
      const double ds = (L.qh + px) + aty;
      double dua_u = fabs(L.dinv * ds);
      double reld = amax(amax(fabs(L.dinv * L.qh), L.dinv * aty), L.dinv * px);
      pri_u = hmax16(pri_u), relp = hmax16(relp), dua_u = hmax16(dua_u), reld = hmax16(reld);
      dua_u *= cinv;
      const double rel_p = relp, rel_d = cinv * reld;
      bool refactor = false;
      if (act) {
        if (pri_u > 1e30 || dua_u > 1e30) {
          status = -7, act = false, iter_done = iter;
        } else if (pri_u < C.eps_abs + C.eps_rel * rel_p && dua_u < C.eps_abs + C.eps_rel * rel_d) {
          status = 1, act = false, iter_done = iter;
        } else if (iter == C.max_iter) {
          status = (pri_u < 10.0 * C.eps_abs + 10.0 * C.eps_rel * rel_p && dua_u < 10.0 * C.eps_abs + 10.0 * C.eps_rel * rel_d) ? 2 : -2;
        }
      }
      if (until_adapt <= 0 && iter % C.adapt_every == 0) {
        until_adapt = C.adapt_every;
        // the scaled norms are needed for the rho estimate only
        double pri_s = amax(fabs(rsA), rsB);
        double np_s = amax(amax(amax(fabs(L.zA), L.zB), axA), axB);
        double dua_s = fabs(ds);
        double nd_s = amax(amax(fabs(L.qh), aty), px);
        pri_s = hmax16(pri_s), np_s = hmax16(np_s), dua_s = hmax16(dua_s), nd_s = hmax16(nd_s);
        if (act) {
          const double pr = pri_s / (np_s + 1e-10);
          const double dr = dua_s / (nd_s + 1e-10);
          double est = rho * sqrt(pr / (dr + 1e-10));
          est = dmin(dmax(est, 1e-6), 1e6);
          if (est > rho * C.adapt_tol || est < rho / C.adapt_tol) {
            rho = est;
            L.rrA = eqA ? 1e3 * rho : rho, L.rrB = eqB ? 1e3 * rho : rho;
            L.riA = 1.0 / L.rrA, L.riB = 1.0 / L.rrB;
            refactor = true;
            ++n_factor;
          }
        }
        if (cta::any(refactor)) factor_pair(C.sigma, C.mu, ps, L.D, L.eA, L.eB, L.rrA, L.rrB, cscale, j, f, c, isv);  // the other half recomputes what it has
      }
      if (!cta::any(act)) break;
    }
  }
  if (act) iter_done = C.max_iter, status = (status == -10) ? -2 : status;
  // ---- retrieve_result (QPWBC.cpp:277-297)
  const bool bad = (status == -7);
  const double df = bad ? nan("") : L.D * L.x;  // delta f
  cta::wsync();
  if (isv) ps[P_V0 + j] = df;
  cta::wsync();
  if (valid && isv) {
    f_res[j] = df + fc;
    st[WS_X + j] = bad ? 0.0 : L.x;
    st[WS_Q + j] = g_;
    st[WS_Z + rA] = bad ? 0.0 : L.zA;
    st[WS_Y + rA] = bad ? 0.0 : L.yA;
    if (hasB) {
      st[WS_Z + rB] = bad ? 0.0 : L.zB;
      st[WS_Y + rB] = bad ? 0.0 : L.yB;
    }
  }
  if (valid && hl < 6) {
    double ar[12], dv[12], s = 0;
    cta::lds12(ps + P_A + 12 * hl, ar);
    cta::lds12(ps + P_V0, dv);
#pragma unroll
    for (int cc = 0; cc < 12; ++cc) s += ar[cc] * dv[cc];
    ddq_res[hl] = s + ps[P_GAM + hl];
  }
  if (valid && hl == 0) {
    st[WS_RHO] = rho;
    info[0] = iter_done;
    info[1] = status;
    info[2] = n_factor;
  }
  cta::wsync();
}

// ------------------------------------------------------------------------------------------
// The leg code: one LEG per lane, lane = 4 * (robot of the group) + leg.
// Layouts: q [19] (xyz, quaternion xyzw, 12 joints), dq [18] (base linear and angular velocity in the base frame, 12
// joints), goals [12] (3 x 4 column-major).
// ------------------------------------------------------------------------------------------
struct BaseState {
  s12::M3 R;
  s12::V3 vb, wb;
};
QRW_HD BaseState load_base(const double* q, const double* dq) {
  BaseState S;
  S.R = s12::quat_to_rot(q[3], q[4], q[5], q[6]);
  S.vb = s12::v3(dq[0], dq[1], dq[2]);
  S.wb = s12::v3(dq[3], dq[4], dq[5]);
  return S;
}
// leg kinematics, InvKin::refreshAndCompute for this foot (Kp = 100, Kd = 20, src/InvKin.hpp:56-57), the foot's rows of the
// contact Jacobian (base columns -> shared memory) and the chain's share of the base wrench of rnea(q, dq, [0; ddq_cmd]).
// `store` = false: a lane without a robot (it computes on the batch's last robot and keeps the shuffles company).
QRW_HDN void pre_leg(const int leg, const bool store, const double* __restrict__ qb, const double* __restrict__ dqb,
                     const double contact_flag, const double* __restrict__ pg, const double* __restrict__ vgp,
                     const double* __restrict__ agp, double* Jc6, double* rnea_s, double* ddq_out, double* vdes_out,
                     double* qdes_out, double* Jc_dbg, double* rnea_dbg) {
  using namespace s12;
  double qj[3], dqj[3];
  for (int c = 0; c < 3; ++c) {
    qj[c] = qb[7 + 3 * leg + c];
    dqj[c] = dqb[6 + 3 * leg + c];
  }
  const LegKin K = leg_kinematics(leg, qj);
  V3 J[3], Ji[3];
  leg_jacobian(K, J);
  V3 vf, wf, acl;
  foot_motion(K, dqj, vf, wf, acl);
  const V3 e = v3(pg[0], pg[1], pg[2]) - K.pf;
  const V3 vg = v3(vgp[0], vgp[1], vgp[2]);
  V3 a = 100.0 * e - 20.0 * (vf - vg) + v3(agp[0], agp[1], agp[2]);
  const bool contact = contact_flag != 0.0;
  if (contact) a = 0.0 * a;
  a = a - acl;  // af + wf x vf of InvKin.cpp:48 = classical acceleration of the foot point
  inv3(J, Ji);
  const V3 dd = a.x * Ji[0] + (a.y * Ji[1] + a.z * Ji[2]);
  const V3 dv = vg.x * Ji[0] + (vg.y * Ji[1] + vg.z * Ji[2]);
  const V3 qs = e.x * Ji[0] + (e.y * Ji[1] + e.z * Ji[2]);
  const double ddq[3] = {dd.x, dd.y, dd.z};
  // rnea(q, dq, [0; ddq_cmd]): base wrench
  const BaseState S = load_base(qb, dqb);
  const V3 ab = cross(S.wb, S.vb) + S12_GRAVITY * v3(S.R.c0.z, S.R.c1.z, S.R.c2.z);  // R' (0, 0, g)
  const LegDyn Dy = leg_dynamics(leg, K, S.wb, v3(0.0, 0.0, 0.0), ab, dqj, ddq);
  double w6[6] = {Dy.f.x, Dy.f.y, Dy.f.z, Dy.n.x, Dy.n.y, Dy.n.z};
#pragma unroll
  for (int c = 0; c < 6; ++c) {  // the four leg lanes of a robot
    w6[c] += cta::hxor(w6[c], 1);
    w6[c] += cta::hxor(w6[c], 2);
  }
  if (leg == 0) {
    V3 fb, nb;
    base_dynamics(S.wb, v3(0.0, 0.0, 0.0), ab, fb, nb);
    const double rn[6] = {w6[0] + fb.x, w6[1] + fb.y, w6[2] + fb.z, w6[3] + nb.x, w6[4] + nb.y, w6[5] + nb.z};
    for (int c = 0; c < 6; ++c) {
      rnea_s[c] = rn[c];
      if (store && rnea_dbg) rnea_dbg[c] = rn[c];
    }
  }
  if (store) {
    for (int c = 0; c < 3; ++c) {
      ddq_out[3 * leg + c] = ddq[c];
      vdes_out[3 * leg + c] = c == 0 ? dv.x : (c == 1 ? dv.y : dv.z);
      qdes_out[3 * leg + c] = qj[c] + (c == 0 ? qs.x : (c == 1 ? qs.y : qs.z));
    }
  }
  // contact Jacobian rows of this foot (LOCAL_WORLD_ALIGNED, linear part): [R | R (e_c x pf) | R J_leg], zero if swinging.
  // The QP reads the six base columns; all 18 go to HBM only when a parity test asks for them.
  const double on = contact ? 1.0 : 0.0;
  for (int c = 0; c < 6; ++c) {
    V3 col;
    if (c < 3)
      col = c == 0 ? S.R.c0 : (c == 1 ? S.R.c1 : S.R.c2);
    else {
      const V3 ec = v3(c == 3 ? 1.0 : 0.0, c == 4 ? 1.0 : 0.0, c == 5 ? 1.0 : 0.0);
      col = mul(S.R, cross(ec, K.pf));
    }
    double* Jr = Jc6 + 3 * leg + 12 * c;
    Jr[0] = on * col.x, Jr[1] = on * col.y, Jr[2] = on * col.z;
    if (store && Jc_dbg) {
      double* Jd = Jc_dbg + 3 * leg + 12 * c;
      Jd[0] = on * col.x, Jd[1] = on * col.y, Jd[2] = on * col.z;
    }
  }
  if (store && Jc_dbg) {
    for (int c = 6; c < 18; ++c) {
      V3 col = v3(0.0, 0.0, 0.0);
      if (c >= 6 + 3 * leg && c < 9 + 3 * leg) col = mul(S.R, J[c - 6 - 3 * leg]);
      double* Jd = Jc_dbg + 3 * leg + 12 * c;
      Jd[0] = on * col.x, Jd[1] = on * col.y, Jd[2] = on * col.z;
    }
  }
}
// tau_ff of the leg: rnea(q, dq, ddq_cmd + [ddq_res; 0])[6:] - Jc[:, 6:]' f_res
QRW_HDN void post_leg(const int leg, const double* __restrict__ qb, const double* __restrict__ dqb, const double contact_flag,
                      const double* ddq_cmd, const double* dr, const double* f_res, double* tau_ff) {
  using namespace s12;
  double qj[3], dqj[3], ddq[3];
  for (int c = 0; c < 3; ++c) {
    qj[c] = qb[7 + 3 * leg + c];
    dqj[c] = dqb[6 + 3 * leg + c];
    ddq[c] = ddq_cmd[3 * leg + c];
  }
  const LegKin K = leg_kinematics(leg, qj);
  V3 J[3];
  leg_jacobian(K, J);
  const BaseState S = load_base(qb, dqb);
  const V3 ab = v3(dr[0], dr[1], dr[2]) + cross(S.wb, S.vb) + S12_GRAVITY * v3(S.R.c0.z, S.R.c1.z, S.R.c2.z);
  const LegDyn Dy = leg_dynamics(leg, K, S.wb, v3(dr[3], dr[4], dr[5]), ab, dqj, ddq);
  const double* fr = f_res + 3 * leg;
  const V3 fl = (contact_flag != 0.0) ? mulT(S.R, v3(fr[0], fr[1], fr[2])) : v3(0.0, 0.0, 0.0);  // R' f
  for (int c = 0; c < 3; ++c) tau_ff[3 * leg + c] = Dy.tau[c] - dot(J[c], fl);
}

// global-memory arguments of a tick
struct TickIo {
  const double *q, *dq, *f_cmd, *contacts, *pgoals, *vgoals, *agoals, *Yinv;
  int f_stride;
  double *st, *f_res, *ddq_res;
  int* info;
  double *ddq_cmd, *vdes, *qdes, *tau_ff, *Jc_dbg, *rnea_dbg;
};

// One GROUP of G = 2, 4 or 8 consecutive robots b0 .. b0 + G - 1 by one warp (every lane calls it with the same arguments):
//   lanes 4r + leg, r < G    leg kinematics, InvKin, contact-Jacobian rows, base wrench -> the warp's slab
//   half-warps, G / 2 rounds  QPWBC of robots b0 + 2p, b0 + 2p + 1
//   lanes 4r + leg, r < G    tau_ff
// A small batch takes small groups (more warps in flight), a large one G = 8 (every lane of the leg code busy).
QRW_HD void tick_group(const WbcConst& C, const int B, const int G, const int first, const TickIo& io, double* sl, const int lane,
                       const int b0) {
  const int r = lane >> 2, leg = lane & 3;
  const bool mine = r < G && b0 + r < B;
  const size_t b = (size_t)(mine ? b0 + r : B - 1);
  const double *qb = io.q + b * 19, *dqb = io.dq + b * 18;
  pre_leg(leg, mine, qb, dqb, io.contacts[b * 4 + leg], io.pgoals + b * 12 + 3 * leg, io.vgoals + b * 12 + 3 * leg,
          io.agoals + b * 12 + 3 * leg, sl + T_JC6 + 72 * r, sl + T_RNEA + 8 * r, io.ddq_cmd + b * 12, io.vdes + b * 12,
          io.qdes + b * 12, io.Jc_dbg ? io.Jc_dbg + b * 216 : nullptr, io.rnea_dbg ? io.rnea_dbg + b * 6 : nullptr);
  cta::wsync();
  const int h = lane >> 4, hl = lane & 15;
#pragma unroll 1
  for (int p = 0; 2 * p < G && b0 + 2 * p < B; ++p) {
    const int rr = 2 * p + h;
    const bool v = b0 + rr < B;
    const size_t bb = (size_t)(v ? b0 + rr : B - 1);
    solve_pair(C, sl + T_HALF + h * P_DOUBLES, hl, v, sl + T_JC6 + 72 * rr, sl + T_RNEA + 8 * rr, io.f_cmd + bb * (size_t)io.f_stride,
               io.st + bb * WBC_STATE_DOUBLES, first, io.Yinv, io.f_res + bb * 12, io.ddq_res + bb * 6, io.info + 4 * bb);
  }
  cta::wsync();  // f_res / ddq_res written by the solver lanes are visible to the leg lanes
  if (mine) post_leg(leg, qb, dqb, io.contacts[b * 4 + leg], io.ddq_cmd + b * 12, io.ddq_res + b * 6, io.f_res + b * 12, io.tau_ff + b * 12);
  cta::wsync();
}

}  // namespace wt
}  // namespace qrw
