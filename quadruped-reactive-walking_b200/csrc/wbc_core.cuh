// wbc_core.cuh -- the whole-body friction-cone QP (QPWBC) and the InvKin step, one instance per
// warp, everything in an 8 KB shared-memory slab.
//
// What it replaces (reference, CPU, one robot):
//   QPWBC::run                         /root/reference/src/QPWBC.cpp:310-343
//     compute_matrices (Y, X, pinv, A, gamma, H, g)     :481-498, include/qrw/InvKin.hpp:60-66
//     update_PQ, bounds l = -G f_cmd, u = l + 25        :520-537, :337-343
//     call_solver -> osqp_setup | osqp_update_P / _lin_cost / _upper_bound / _lower_bound, osqp_solve  :213-275
//     retrieve_result                                   :277-297
//   InvKin::refreshAndCompute          /root/reference/src/InvKin.cpp:23-73
// The 12-variable / 20-row QP is solved by the same OSQP-equivalent ADMM as the MPC (Ruiz x10
// recomputed every call from the new P and the PREVIOUS call's q -- osqp_update_P rescales before
// osqp_update_lin_cost replaces q --, rho = 0.1 with adaptation every 200 iterations, eps =
// (float)1e-5, warm start).  The KKT system is reduced to the dense SPD 12 x 12
// K = P + sigma I + A' diag(rho) A and inverted explicitly once per factorisation.
//
// Infeasibility exits are not implemented: the QP is always strictly feasible (f = (0,0,eps) per
// foot lies inside 0 <= G f <= 25) and P >= 5 I, so OSQP's certificates cannot be produced.
#pragma once
#include "qrw_spmd.cuh"

namespace qrw {

// slab layout (doubles)
constexpr int W_A = 0;        // 6 x 12  A = Yinv X (row-major)
constexpr int W_GAM = 72;     // 6
constexpr int W_H = 80;       // 12 x 12 (row-major, symmetric)
constexpr int W_G = 224;      // 12   linear cost g
constexpr int W_D = 236;      // 12
constexpr int W_E = 248;      // 20
constexpr int W_KINV = 268;   // 12 x 12
constexpr int W_X = 412;      // 12
constexpr int W_Z = 424;      // 20
constexpr int W_Y = 444;      // 20
constexpr int W_XT = 464;     // 12
constexpr int W_RHS = 476;    // 12
constexpr int W_L = 488;      // 20 scaled bounds
constexpr int W_U = 508;      // 20
constexpr int W_Q = 528;      // 12 scaled q
constexpr int W_EV = 540;     // 20
constexpr int W_QOLD = 560;   // 12
constexpr int W_YINV = 572;   // 36 (column-major)
constexpr int W_TMP = 608;    // 384 scratch (Jacobi U,V / Gauss-Jordan / reductions)
constexpr int WBC_SLAB_DOUBLES = 992;
constexpr int WBC_SLAB_BYTES = WBC_SLAB_DOUBLES * 8;

// persistent state per instance: x[12] z[20] y[20] q_prev[12] rho  -> 65, padded
constexpr int WS_X = 0, WS_Z = 12, WS_Y = 32, WS_Q = 52, WS_RHO = 64;
constexpr int WBC_STATE_DOUBLES = 72;

struct WbcConst {
  double mu, sigma, alpha, eps_abs, eps_rel, rho0, adapt_tol, q1, q2, fmax;
  int max_iter, check_every, adapt_every;
};
inline WbcConst make_wbc_const() {
  WbcConst C;
  C.mu = 0.9;  // double literal in QPWBC.hpp:30 (NOT the MPC's 0.9f)
  C.sigma = 1e-6;
  C.alpha = 1.6;
  C.eps_abs = (float)1e-5;  // QPWBC.cpp:239-240
  C.eps_rel = (float)1e-5;
  C.rho0 = 0.1;
  C.adapt_tol = 5.0;
  C.q1 = 0.1;  // Q1 = 0.1 I6, Q2 = 5 I12 (QPWBC.hpp:26-27)
  C.q2 = 5.0;
  C.fmax = 25.0;
  C.max_iter = 4000;
  C.check_every = 25;
  C.adapt_every = 200;
  return C;
}

// G (20 x 12): per foot [[-1,0,mu],[1,0,mu],[0,-1,mu],[0,1,mu],[0,0,1]]  (QPWBC.cpp:4-22)
// (G v)[r] for a 12-vector v
QRW_HD double wbc_g_row(double mu, const double* v, int r) {
  const int f = r / 5, c = r - 5 * f;
  const double vz = v[3 * f + 2];
  if (c == 4) return vz;
  const double vt = v[3 * f + (c >> 1)];
  return ((c & 1) ? vt : -vt) + mu * vz;
}
// (G' w)[j] for a 20-vector w
QRW_HD double wbc_gt_col(double mu, const double* w, int j) {
  const int f = j / 3, c = j - 3 * f;
  const double* wf = w + 5 * f;
  if (c == 0) return wf[1] - wf[0];
  if (c == 1) return wf[3] - wf[2];
  return mu * (wf[0] + wf[1] + wf[2] + wf[3]) + wf[4];
}
QRW_HD double wbc_g_entry(double mu, int r, int j) {
  const int f = r / 5, c = r - 5 * f;
  if (j / 3 != f) return 0.0;
  const int cj = j - 3 * f;
  if (c == 4) return cj == 2 ? 1.0 : 0.0;
  if (cj == 2) return mu;
  if (cj == (c >> 1)) return (c & 1) ? 1.0 : -1.0;
  return 0.0;
}

// pseudoInverse of a column-major 6 x 6 (include/qrw/InvKin.hpp:60-66), one-sided Jacobi SVD,
// run by a single lane (Y is diagonal in practice -- scripts/QP_WBC.py:89-93 -- so the sweep
// finds nothing to rotate).  U, V: 36 doubles each of scratch.
QRW_HD void pinv6_serial(const double* a, double* out, double* U, double* V) {
  for (int c = 0; c < 6; ++c)
    for (int r = 0; r < 6; ++r) {
      U[6 * c + r] = a[r + 6 * c];
      V[6 * c + r] = (r == c) ? 1.0 : 0.0;
    }
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < 5; ++p)
      for (int q = p + 1; q < 6; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int r = 0; r < 6; ++r) {
          alpha += U[6 * p + r] * U[6 * p + r];
          beta += U[6 * q + r] * U[6 * q + r];
          gamma += U[6 * p + r] * U[6 * q + r];
        }
        if (gamma == 0.0 || fabs(gamma) <= 1e-15 * sqrt(alpha * beta)) continue;
        rotated = true;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = ((zeta >= 0) ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
        for (int r = 0; r < 6; ++r) {
          const double up = U[6 * p + r], uq = U[6 * q + r];
          U[6 * p + r] = cs * up - sn * uq;
          U[6 * q + r] = sn * up + cs * uq;
          const double vp = V[6 * p + r], vq = V[6 * q + r];
          V[6 * p + r] = cs * vp - sn * vq;
          V[6 * q + r] = sn * vp + cs * vq;
        }
      }
    if (!rotated) break;
  }
  double sv[6], svmax = 0.0;
  for (int c = 0; c < 6; ++c) {
    double s = 0;
    for (int r = 0; r < 6; ++r) s += U[6 * c + r] * U[6 * c + r];
    sv[c] = sqrt(s);
    svmax = dmax(svmax, sv[c]);
  }
  const double tol = 2.220446049250313e-16 * 6.0 * svmax;
  for (int i = 0; i < 36; ++i) out[i] = 0.0;
  for (int c = 0; c < 6; ++c) {
    if (!(sv[c] > tol)) continue;
    const double inv2 = 1.0 / (sv[c] * sv[c]);
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) out[i + 6 * j] += V[6 * c + i] * inv2 * U[6 * c + j];
  }
}

// P as OSQP stores it: the upper triangle of H, mirrored (update_PQ, QPWBC.cpp:520-532)
QRW_HD double wbc_hs(const double* H, int i, int j) { return i <= j ? H[12 * i + j] : H[12 * j + i]; }

QRW_HD bool wbc_row_eq(const double* l, const double* u, int r) { return (u[r] - l[r]) < 1e-4; }

// K = P_hat + sigma I + A_hat' diag(rho) A_hat, inverted in place (Gauss-Jordan, SPD)
QRW_HD void wbc_factor(const WbcConst& C, double* sl, double rho, double cscale) {
  double* K = sl + W_KINV;
  double* tp = sl + W_TMP;
  const double *H = sl + W_H, *D = sl + W_D, *E = sl + W_E, *lo = sl + W_L, *up = sl + W_U;
  QRW_LANES(l) {
    for (int e = l; e < 144; e += 32) {
      const int i = e / 12, j = e - 12 * i;
      double v = cscale * D[i] * wbc_hs(H, i, j) * D[j];
      if (i == j) v += C.sigma;
      if (i / 3 == j / 3) {
        const int f = i / 3;
        double acc = 0.0;
        for (int c = 0; c < 5; ++c) {
          const int r = 5 * f + c;
          const double rr = wbc_row_eq(lo, up, r) ? 1e3 * rho : rho;
          acc += rr * (E[r] * wbc_g_entry(C.mu, r, i)) * (E[r] * wbc_g_entry(C.mu, r, j));
        }
        v += acc * D[i] * D[j];
      }
      K[e] = v;
    }
  }
  QRW_SYNC();
  for (int p = 0; p < 12; ++p) {
    const double d = 1.0 / K[12 * p + p];
    QRW_LANES(l) {
      if (l < 12) {
        tp[l] = (l == p) ? d : K[12 * p + l] * d;
        tp[12 + l] = K[12 * l + p];
      }
    }
    QRW_SYNC();
    QRW_LANES(l) {
      for (int e = l; e < 144; e += 32) {
        const int r = e / 12, c2 = e - 12 * r;
        double v;
        if (r == p)
          v = tp[c2];
        else if (c2 == p)
          v = -tp[12 + r] * d;
        else
          v = K[e] - tp[12 + r] * tp[c2];
        K[e] = v;
      }
    }
    QRW_SYNC();
  }
}

// One QPWBC::run.  M 18x18, Jc 12x18 column-major.  st: persistent state.  Outputs f_res[12],
// ddq_res[6], optional H[144] (column-major), info {iters, status, n_factor}.
// Yinv_pre (optional): pinv(M[:6, :6]) computed beforehand by the same pinv6_serial -- the whole-body tick passes one M for
// the whole batch (crba at the zero configuration is a model constant), so its pseudo-inverse is computed once per handle.
QRW_HD void wbc_solve_instance(const WbcConst& C, double* sl, const double* M, const double* Jc,
                               const double* f_cmd, const double* RNEA, double* st, int first,
                               double* f_res, double* ddq_res, double* H_out, int* info, const double* Yinv_pre = nullptr) {
  double *A = sl + W_A, *gam = sl + W_GAM, *H = sl + W_H, *g = sl + W_G, *D = sl + W_D, *E = sl + W_E,
         *Kinv = sl + W_KINV, *x = sl + W_X, *z = sl + W_Z, *y = sl + W_Y, *xt = sl + W_XT,
         *rhs = sl + W_RHS, *lo = sl + W_L, *up = sl + W_U, *qh = sl + W_Q, *ev = sl + W_EV,
         *qold = sl + W_QOLD, *Yinv = sl + W_YINV, *tmp = sl + W_TMP;
  // ---- compute_matrices (QPWBC.cpp:481-498)
  QRW_LANES(l) {
    for (int e = l; e < 36; e += 32) tmp[72 + e] = M[(e % 6) + 18 * (e / 6)];  // Y, column-major
    if (l < 12) x[l] = first ? 0.0 : st[WS_X + l];
    if (l < 20) {
      z[l] = first ? 0.0 : st[WS_Z + l];
      y[l] = first ? 0.0 : st[WS_Y + l];
    }
  }
  QRW_SYNC();
  if (Yinv_pre) {
    QRW_LANES(l) {
      for (int e = l; e < 36; e += 32) Yinv[e] = Yinv_pre[e];
    }
  } else {
    QRW_LANES(l) {
      if (l == 0) pinv6_serial(tmp + 72, Yinv, tmp, tmp + 36);
    }
  }
  QRW_SYNC();
  double* Xf = tmp;  // X f_cmd - RNEA (6)
  QRW_LANES(l) {
    for (int e = l; e < 72; e += 32) {  // A = Yinv X, X[r][c] = Jc(c, r)
      const int r = e / 12, c = e - 12 * r;
      double s = 0;
      for (int k = 0; k < 6; ++k) s += Yinv[r + 6 * k] * Jc[c + 12 * k];
      A[e] = s;
    }
    if (l < 6) {
      double s = 0;
      for (int c = 0; c < 12; ++c) s += Jc[c + 12 * l] * f_cmd[c];
      Xf[l] = s - RNEA[l];
    }
  }
  QRW_SYNC();
  QRW_LANES(l) {
    if (l < 6) {
      double s = 0;
      for (int k = 0; k < 6; ++k) s += Yinv[l + 6 * k] * Xf[k];
      gam[l] = s;
    }
  }
  QRW_SYNC();
  QRW_LANES(l) {
    for (int e = l; e < 144; e += 32) {  // H = A' Q1 A + Q2
      const int i = e / 12, j = e - 12 * i;
      double s = 0;
      for (int k = 0; k < 6; ++k) s += (A[12 * k + i] * C.q1) * A[12 * k + j];
      H[e] = s + (i == j ? C.q2 : 0.0);
      if (H_out) H_out[i + 12 * j] = H[e];
    }
    if (l < 12) {
      double s = 0;
      for (int k = 0; k < 6; ++k) s += (A[12 * k + l] * C.q1) * gam[k];
      g[l] = s;
      // osqp_update_P rescales with the q still held by the workspace (previous call's g)
      qold[l] = first ? s : st[WS_Q + l];
    }
    if (l < 20) {  // bounds (QPWBC.cpp:337-343), unscaled for now
      const double Gf = wbc_g_row(C.mu, f_cmd, l);
      lo[l] = -Gf;
      up[l] = -Gf + C.fmax;
    }
  }
  QRW_SYNC();
  // ---- Ruiz equilibration (OSQP scale_data) on P = H (dense symmetric), A = G, q = qold
  double cscale = 1.0;
  QRW_LANES(l) {
    if (l < 12) D[l] = 1.0;
    if (l < 20) E[l] = 1.0;
  }
  QRW_SYNC();
  for (int pass = 0; pass < 10; ++pass) {
    QRW_LANES(l) {
      if (l < 12) {
        const int j = l;
        double nrm = 0.0;
        for (int i = 0; i < 12; ++i) nrm = dmax(nrm, fabs(cscale * D[i] * wbc_hs(H, i, j) * D[j]));
        for (int r = 5 * (j / 3); r < 5 * (j / 3) + 5; ++r) nrm = dmax(nrm, fabs(E[r] * wbc_g_entry(C.mu, r, j) * D[j]));
        tmp[l] = 1.0 / sqrt(limit_scaling(nrm));
      }
      if (l < 20) {
        const int f = l / 5;
        double nrm = 0.0;
        for (int c = 0; c < 3; ++c) nrm = dmax(nrm, fabs(E[l] * wbc_g_entry(C.mu, l, 3 * f + c) * D[3 * f + c]));
        tmp[32 + l] = 1.0 / sqrt(limit_scaling(nrm));
      }
    }
    QRW_SYNC();
    QRW_LANES(l) {
      if (l < 12) D[l] *= tmp[l];
      if (l < 20) E[l] *= tmp[32 + l];
    }
    QRW_SYNC();
    QRW_LANES(l) {
      if (l < 12) {
        double nrm = 0.0;
        for (int i = 0; i < 12; ++i) nrm = dmax(nrm, fabs(cscale * D[i] * wbc_hs(H, i, l) * D[l]));
        tmp[64 + l] = nrm;
        tmp[96 + l] = fabs(cscale * D[l] * qold[l]);
      }
    }
    QRW_SYNC();
    double mean = 0.0, qn = 0.0;
    for (int j = 0; j < 12; ++j) {
      mean += tmp[64 + j];
      qn = dmax(qn, tmp[96 + j]);
    }
    mean /= 12.0;
    qn = limit_scaling(qn);
    double c_temp = limit_scaling(dmax(mean, qn));
    cscale *= 1.0 / c_temp;
    QRW_SYNC();
  }
  const double cinv = 1.0 / cscale;
  QRW_LANES(l) {
    if (l < 12) qh[l] = cscale * (D[l] * g[l]);  // osqp_update_lin_cost
    if (l < 20) {
      lo[l] = E[l] * lo[l];
      up[l] = E[l] * up[l];
    }
  }
  QRW_SYNC();
  double rho = first ? C.rho0 : st[WS_RHO];
  rho = dmin(dmax(rho, 1e-6), 1e6);
  int n_factor = 1;
  wbc_factor(C, sl, rho, cscale);

  // ---- ADMM
  int status = -10, iter = 0;
  for (iter = 1; iter <= C.max_iter; ++iter) {
    QRW_LANES(l) {
      if (l < 20) {
        const double rr = wbc_row_eq(lo, up, l) ? 1e3 * rho : rho;
        ev[l] = E[l] * (rr * z[l] - y[l]);
      }
    }
    QRW_SYNC();
    QRW_LANES(l) {
      if (l < 12) rhs[l] = C.sigma * x[l] - qh[l] + D[l] * wbc_gt_col(C.mu, ev, l);
    }
    QRW_SYNC();
    QRW_LANES(l) {
      if (l < 12) {
        double s0 = 0, s1 = 0;
        for (int i = 0; i < 12; i += 2) {
          s0 += Kinv[12 * l + i] * rhs[i];
          s1 += Kinv[12 * l + i + 1] * rhs[i + 1];
        }
        xt[l] = s0 + s1;
      }
    }
    QRW_SYNC();
    QRW_LANES(l) {
      if (l < 12) ev[l] = D[l] * xt[l];  // ev reused as D .* x_tilde (12)
    }
    QRW_SYNC();
    QRW_LANES(l) {
      if (l < 20) {
        const bool eq = wbc_row_eq(lo, up, l);
        const double rr = eq ? 1e3 * rho : rho, ri = 1.0 / rr;
        const double zt = E[l] * wbc_g_row(C.mu, ev, l);
        const double zr = C.alpha * zt + (1.0 - C.alpha) * z[l];
        const double zn = dmin(dmax(zr + ri * y[l], lo[l]), up[l]);
        y[l] += rr * (zr - zn);
        z[l] = zn;
      }
      if (l < 12) x[l] = C.alpha * xt[l] + (1.0 - C.alpha) * x[l];
    }
    QRW_SYNC();
    if (iter % C.check_every == 0) {
      QRW_LANES(l) {
        if (l < 12) tmp[l] = D[l] * x[l];
        if (l < 20) tmp[32 + l] = E[l] * y[l];
      }
      QRW_SYNC();
      QRW_LANES(l) {
        if (l < 20) {
          const double ax = E[l] * wbc_g_row(C.mu, tmp, l);
          tmp[64 + l] = ax;
        }
        if (l < 12) {
          double px = 0.0;
          for (int i = 0; i < 12; ++i) px += (cscale * D[i] * wbc_hs(H, i, l) * D[l]) * x[i];
          tmp[96 + l] = px;
          tmp[128 + l] = D[l] * wbc_gt_col(C.mu, tmp + 32, l);
        }
      }
      QRW_SYNC();
      double pri_s = 0, pri_u = 0, nz_s = 0, nz_u = 0, nax_s = 0, nax_u = 0;
      for (int r = 0; r < 20; ++r) {
        const double ax = tmp[64 + r], einv = 1.0 / E[r], rs = ax - z[r];
        pri_s = dmax(pri_s, fabs(rs));
        pri_u = dmax(pri_u, fabs(einv * rs));
        nz_s = dmax(nz_s, fabs(z[r]));
        nz_u = dmax(nz_u, fabs(einv * z[r]));
        nax_s = dmax(nax_s, fabs(ax));
        nax_u = dmax(nax_u, fabs(einv * ax));
      }
      double dua_s = 0, dua_u = 0, nq_s = 0, nq_u = 0, naty_s = 0, naty_u = 0, npx_s = 0, npx_u = 0;
      for (int j = 0; j < 12; ++j) {
        const double px = tmp[96 + j], aty = tmp[128 + j], dinv = 1.0 / D[j];
        const double ds = (qh[j] + px) + aty;
        dua_s = dmax(dua_s, fabs(ds));
        dua_u = dmax(dua_u, fabs(dinv * ds));
        nq_s = dmax(nq_s, fabs(qh[j]));
        nq_u = dmax(nq_u, fabs(dinv * qh[j]));
        naty_s = dmax(naty_s, fabs(aty));
        naty_u = dmax(naty_u, fabs(dinv * aty));
        npx_s = dmax(npx_s, fabs(px));
        npx_u = dmax(npx_u, fabs(dinv * px));
      }
      dua_u *= cinv;
      QRW_SYNC();
      if (pri_u > 1e30 || dua_u > 1e30) {
        status = -7;
        break;
      }
      const double rel_p = dmax(nz_u, nax_u), rel_d = cinv * dmax(dmax(nq_u, naty_u), npx_u);
      if (pri_u < C.eps_abs + C.eps_rel * rel_p && dua_u < C.eps_abs + C.eps_rel * rel_d) {
        status = 1;
        break;
      }
      if (iter == C.max_iter)
        status = (pri_u < 10.0 * C.eps_abs + 10.0 * C.eps_rel * rel_p && dua_u < 10.0 * C.eps_abs + 10.0 * C.eps_rel * rel_d) ? 2 : -2;
      if (iter % C.adapt_every == 0) {
        const double pr = pri_s / (dmax(nz_s, nax_s) + 1e-10);
        const double dr = dua_s / (dmax(dmax(nq_s, naty_s), npx_s) + 1e-10);
        double est = rho * sqrt(pr / (dr + 1e-10));
        est = dmin(dmax(est, 1e-6), 1e6);
        if (est > rho * C.adapt_tol || est < rho / C.adapt_tol) {
          rho = est;
          wbc_factor(C, sl, rho, cscale);
          ++n_factor;
        }
      }
    }
  }
  if (iter > C.max_iter) iter = C.max_iter;
  if (status == -10) status = -2;
  // ---- retrieve_result (QPWBC.cpp:277-297)
  const bool bad = (status == -7);
  QRW_LANES(l) {
    if (l < 12) ev[l] = bad ? nan("") : D[l] * x[l];  // delta f
  }
  QRW_SYNC();
  QRW_LANES(l) {
    if (l < 12) {
      f_res[l] = ev[l] + f_cmd[l];
      st[WS_X + l] = bad ? 0.0 : x[l];
      st[WS_Q + l] = g[l];
    }
    if (l < 6) {
      double s = 0;
      for (int c = 0; c < 12; ++c) s += A[12 * l + c] * ev[c];
      ddq_res[l] = s + gam[l];
    }
    if (l < 20) {
      st[WS_Z + l] = bad ? 0.0 : z[l];
      st[WS_Y + l] = bad ? 0.0 : y[l];
    }
    if (l == 0) {
      st[WS_RHO] = rho;
      info[0] = iter;
      info[1] = status;
      info[2] = n_factor;
    }
  }
  QRW_SYNC();
}

// ------------------------------------------------------------------------------------------
// InvKin::refreshAndCompute (src/InvKin.cpp:23-73) for one instance; lanes 0..3 = feet.
// ------------------------------------------------------------------------------------------
QRW_HD void invkin_instance(const double* contacts, const double* goals, const double* vgoals,
                            const double* agoals, const double* posf, const double* vf, const double* wf,
                            const double* af, const double* Jf, double* ddq, double* dq_cmd, double* q_step) {
  const double Kp = 100.0, Kd = 20.0;  // Kd = 2 sqrt(Kp), InvKin.hpp:56-57
  QRW_LANES(l) {
    if (l < 4) {
      const int i = l;
      double e[3], a[3], v[3], w[3], vg[3];
      for (int c = 0; c < 3; ++c) {
        v[c] = vf[i + 4 * c];
        w[c] = wf[i + 4 * c];
        vg[c] = vgoals[c + 3 * i];
        e[c] = goals[c + 3 * i] - posf[i + 4 * c];
        a[c] = Kp * e[c] - Kd * (v[c] - vg[c]) + agoals[c + 3 * i];
      }
      if (contacts[i] != 0.0)
        for (int c = 0; c < 3; ++c) a[c] *= 0.0;
      const double cr[3] = {w[1] * v[2] - w[2] * v[1], w[2] * v[0] - w[0] * v[2], w[0] * v[1] - w[1] * v[0]};
      for (int c = 0; c < 3; ++c) a[c] -= af[i + 4 * c] + cr[c];
      double Mm[3][3], I[3][3];
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) Mm[r][c] = Jf[(3 * i + r) + 12 * (3 * i + c)];
      const double c00 = Mm[1][1] * Mm[2][2] - Mm[1][2] * Mm[2][1];
      const double c10 = Mm[1][2] * Mm[2][0] - Mm[1][0] * Mm[2][2];
      const double c20 = Mm[1][0] * Mm[2][1] - Mm[1][1] * Mm[2][0];
      const double det = c00 * Mm[0][0] + c10 * Mm[0][1] + c20 * Mm[0][2];
      const double id = 1.0 / det;
      I[0][0] = c00 * id;
      I[1][0] = c10 * id;
      I[2][0] = c20 * id;
      I[0][1] = (Mm[0][2] * Mm[2][1] - Mm[0][1] * Mm[2][2]) * id;
      I[1][1] = (Mm[0][0] * Mm[2][2] - Mm[0][2] * Mm[2][0]) * id;
      I[2][1] = (Mm[0][1] * Mm[2][0] - Mm[0][0] * Mm[2][1]) * id;
      I[0][2] = (Mm[0][1] * Mm[1][2] - Mm[0][2] * Mm[1][1]) * id;
      I[1][2] = (Mm[0][2] * Mm[1][0] - Mm[0][0] * Mm[1][2]) * id;
      I[2][2] = (Mm[0][0] * Mm[1][1] - Mm[0][1] * Mm[1][0]) * id;
      for (int r = 0; r < 3; ++r) {
        double s0 = 0, s1 = 0, s2 = 0;
        for (int c = 0; c < 3; ++c) {
          s0 += I[r][c] * a[c];
          s1 += I[r][c] * vg[c];
          s2 += I[r][c] * e[c];
        }
        ddq[3 * i + r] = s0;
        dq_cmd[3 * i + r] = s1;
        q_step[3 * i + r] = s2;
      }
    }
  }
}

}  // namespace qrw
