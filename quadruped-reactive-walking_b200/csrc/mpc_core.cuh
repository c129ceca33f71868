// mpc_core.cuh -- one Solo12 centroidal-MPC solve, executed by ONE warp out of a 75 KB
// shared-memory slab (three resident warps per B200 SM).
//
// What it replaces (reference, CPU, one robot):
//   MPC::run                      /root/reference/src/MPC.cpp:626-649
//     construct_gait / construct_S            :686-701, :665-681
//     create_ML / update_ML (B_k, S flags)    :74-256, :418-464
//     create_NK / update_NK (bounds)          :261-312, :469-496
//     call_solver -> osqp_setup | osqp_update_A + osqp_update_bounds, osqp_solve   :501-564
//     retrieve_result                         :569-599
// OSQP (v0.6.x, not vendored in the reference) is replaced by an ADMM that follows OSQP's
// iterate sequence exactly -- same Ruiz equilibration (10 passes, recomputed every call), same
// rho vector (1e3*rho on equality rows), sigma, alpha, termination test every 25 iterations,
// rho adaptation every 200, warm start from the previous call's scaled iterates -- and differs
// only in how the KKT system is solved.
//
// Linear algebra (B200-first): the quasi-definite KKT system is reduced to
//     K = P + sigma I + A' diag(rho) A            (384 x 384, SPD)
// and, with the variables ordered by time [f_0 X_1 | f_1 X_2 | ... | f_15 X_16], K is block
// tridiagonal with 16 blocks of 24.  It is factored as block LDL' with explicitly inverted 24x24
// pivot blocks S_k^-1 kept packed-symmetric in shared memory (16 * 300 doubles = 38.4 KB); the
// sub-diagonal blocks are never stored: they are rank-12 products of the (structurally known)
// constraint rows and are applied on the fly.  One ADMM iteration = forward sweep (16 dense
// 24x24 symmetric mat-vecs) + backward sweep (15 dense 24x12) + two structured passes over the
// 1998-nnz constraint matrix, all fp64 out of shared memory, lane = row.
//
// The same source compiles as a host-side warp emulation (see qrw_spmd.cuh) for CPU tests.
#pragma once
#include "mpc_common.cuh"

namespace qrw {

// shared slab layout (doubles)
constexpr int OFF_SINV = 0;                 // 16 * 300 packed symmetric pivot-block inverses
constexpr int OFF_X = OFF_SINV + NS * 300;  // x   (scaled iterate)
constexpr int OFF_Z = OFF_X + NX;           // z
constexpr int OFF_Y = OFF_Z + NR;           // y
constexpr int OFF_D = OFF_Y + NR;           // Ruiz column scaling
constexpr int OFF_E = OFF_D + NX;           // Ruiz row scaling
constexpr int OFF_BND = OFF_E + NR;         // scaled equality bounds of the 192 dynamics rows
constexpr int OFF_BANG = OFF_BND + 192;     // 16 * (3 x 12) angular blocks of B_k (unscaled)
constexpr int OFF_WK = OFF_BANG + 576;      // work: rhs | x_tilde | small temporaries
constexpr int WK_SIZE = 896;
constexpr int OFF_INT = OFF_WK + WK_SIZE;   // 128 ints
constexpr int MPC_SLAB_DOUBLES = OFF_INT + 64;
constexpr int MPC_SLAB_BYTES = MPC_SLAB_DOUBLES * 8;  // 75 264 B -> 3 warps per SM

struct MpcSlab {
  double *Sinv, *x, *z, *y, *D, *E, *bnd, *Bang, *wk;
  int* ib;
  QRW_HD explicit MpcSlab(double* base)
      : Sinv(base + OFF_SINV), x(base + OFF_X), z(base + OFF_Z), y(base + OFF_Y), D(base + OFF_D),
        E(base + OFF_E), bnd(base + OFF_BND), Bang(base + OFF_BANG), wk(base + OFF_WK),
        ib(reinterpret_cast<int*>(base + OFF_INT)) {}
};

QRW_HD int sym_idx(int a, int b) { return a >= b ? (a * (a + 1)) / 2 + b : (b * (b + 1)) / 2 + a; }
// ------------------------------------------------------------------------------------------
// Structured products with the UNSCALED constraint matrix A0 (pattern of ML, MPC.cpp:74-146).
// Rows: [0,192) dynamics (k,i); [192,384) force-enable (k,j); [384,704) friction cone (k,foot,c).
// ------------------------------------------------------------------------------------------
// (A0 * (D .* v))[row] for a dynamics row
QRW_HD double a0_row_dyn(const MpcConst& C, const double* D, const double* Bang, const double* v, int k, int i) {
  double s = -(D[12 * k + i] * v[12 * k + i]);
  if (k > 0) {
    const int p = 12 * (k - 1) + i;
    s += D[p] * v[p];
    if (i < 6) s += C.dt * (D[p + 6] * v[p + 6]);
  }
  if (i >= 6) {
    const double* Df = D + 192 + 12 * k;
    const double* vf = v + 192 + 12 * k;
    if (i < 9) {
      const int c = i - 6;
      s += C.dtm * (Df[c] * vf[c] + Df[3 + c] * vf[3 + c] + Df[6 + c] * vf[6 + c] + Df[9 + c] * vf[9 + c]);
    } else {
      const double* Br = Bang + 36 * k + 12 * (i - 9);
      double t = 0.0;
#pragma unroll
      for (int a = 0; a < 12; ++a) t += Br[a] * (Df[a] * vf[a]);
      s += t;
    }
  }
  return s;
}
// (A0 * (D .* v))[row] for a friction-cone row q = 20k + 5 foot + c
QRW_HD double a0_row_cone(const MpcConst& C, const double* D, const double* v, int q) {
  const int k = q / 20, rem = q - 20 * k, f = rem / 5, c = rem - 5 * f;
  const int b = 192 + 12 * k + 3 * f;
  const double fz = D[b + 2] * v[b + 2];
  if (c == 4) return -fz;
  const int ax = b + (c >> 1);
  const double ft = D[ax] * v[ax];
  return ((c & 1) ? -ft : ft) - C.mu * fz;
}
// (A0' * u)[var] with u(row) supplied by a functor; state variable X_{k+1}[i]
template <class U>
QRW_HD double a0t_col_state(const MpcConst& C, int k, int i, U u) {
  double s = -u(12 * k + i);
  if (k < NS - 1) {
    s += u(12 * (k + 1) + i);
    if (i >= 6) s += C.dt * u(12 * (k + 1) + i - 6);
  }
  return s;
}
// force variable f_k[a]
template <class U>
QRW_HD double a0t_col_force(const MpcConst& C, const double* Bang, const int* sb, int k, int a, U u) {
  const int f = a / 3, c = a - 3 * f;
  double s = C.dtm * u(12 * k + 6 + c);
  s += Bang[36 * k + a] * u(12 * k + 9);
  s += Bang[36 * k + 12 + a] * u(12 * k + 10);
  s += Bang[36 * k + 24 + a] * u(12 * k + 11);
  if ((sb[k] >> f) & 1) s += u(192 + 12 * k + a);
  const int r0 = 384 + 20 * k + 5 * f;
  if (c == 0)
    s += u(r0) - u(r0 + 1);
  else if (c == 1)
    s += u(r0 + 2) - u(r0 + 3);
  else
    s += -C.mu * (u(r0) + u(r0 + 1) + u(r0 + 2) + u(r0 + 3)) - u(r0 + 4);
  return s;
}

// scaled dynamics-row entry G[i][a] = E * A0 * D of row (k,i) on stage-k variable a
QRW_HD double g_entry(const MpcConst& C, const MpcSlab& S, int k, int i, int a) {
  if (a >= 12) return (a - 12 == i) ? -(S.E[12 * k + i] * S.D[12 * k + i]) : 0.0;
  if (i < 6) return 0.0;
  const double ed = S.E[12 * k + i] * S.D[192 + 12 * k + a];
  if (i < 9) return (a % 3 == i - 6) ? C.dtm * ed : 0.0;
  return S.Bang[36 * k + 12 * (i - 9) + a] * ed;
}
// the (up to 4) dynamics rows of stage k that touch stage variable a
QRW_HD int g_rows(int a, int rows[4]) {
  if (a >= 12) {
    rows[0] = a - 12;
    return 1;
  }
  rows[0] = 6 + a % 3;
  rows[1] = 9;
  rows[2] = 10;
  rows[3] = 11;
  return 4;
}

// rho of a constraint row: equality rows (u - l < RHO_TOL in scaled bounds) get 1e3 * rho.
// Cone rows have l = -inf except every 5th, whose scaled width is 25 * E.
QRW_HD bool row_is_eq(const MpcSlab& S, int r) {
  if (r < 384) return true;
  const int q = r - 384;
  return (q % 5 == 4) && (25.0 * S.E[r] < 1e-4);
}

// ------------------------------------------------------------------------------------------
// Block LDL' of K: pivot-block inverses into S.Sinv (packed symmetric).
// ------------------------------------------------------------------------------------------
QRW_HD void mpc_factor(const MpcConst& C, const MpcSlab& S, const RhoSet R, double cscale) {
  double* W = S.wk;         // 24 x 24 pivot block (row-major)
  double* N = S.wk + 576;   // 12 x 12:  Jx * Sinv_{k-1}[x,x] * Jx'
  double* tp = S.wk + 720;  // 48 temporaries for the Gauss-Jordan sweep
  const int* sb = S.ib + 32;
  for (int k = 0; k < NS; ++k) {
    if (k > 0) {
      const double* Mp = S.Sinv + 300 * (k - 1);
      QRW_LANES(l) {
        for (int e = l; e < 144; e += 32) {
          const int i = e / 12, j = e - 12 * i;
          // Jx row i: columns i (E D) and i+6 (dt E D), i < 6
          const double ei = S.E[12 * k + i], ej = S.E[12 * k + j];
          const double* Dp = S.D + 12 * (k - 1);
          double acc = 0.0;
          const int ni = (i < 6) ? 2 : 1, nj = (j < 6) ? 2 : 1;
          for (int s = 0; s < ni; ++s) {
            const int p = i + 6 * s;
            const double cp = (s ? C.dt : 1.0) * ei * Dp[p];
            for (int t = 0; t < nj; ++t) {
              const int q = j + 6 * t;
              const double cq = (t ? C.dt : 1.0) * ej * Dp[q];
              acc += cp * Mp[sym_idx(12 + p, 12 + q)] * cq;
            }
          }
          N[e] = acc;
        }
      }
      QRW_SYNC();
    }
    QRW_LANES(l) {
      for (int e = l; e < 576; e += 32) {
        const int a = e / 24, b = e - 24 * a;
        int ra[4], rb[4];
        const int na = g_rows(a, ra), nb = g_rows(b, rb);
        double ga[4], gb[4];
        for (int s = 0; s < na; ++s) ga[s] = g_entry(C, S, k, ra[s], a);
        for (int t = 0; t < nb; ++t) gb[t] = g_entry(C, S, k, rb[t], b);
        double val = 0.0;
        // rho_eq * (dynamics rows of stage k)' (dynamics rows of stage k)
        for (int s = 0; s < na; ++s)
          for (int t = 0; t < nb; ++t)
            if (ra[s] == rb[t]) val += ga[s] * gb[t];
        val *= R.rho_eq;
        if (a == b) {
          if (a < 12) {
            const double d = S.D[192 + 12 * k + a];
            val += cscale * C.wf * d * d + C.sigma;
            if ((sb[k] >> (a / 3)) & 1) {
              const double sd = S.E[192 + 12 * k + a] * d;
              val += R.rho_eq * sd * sd;
            }
          } else {
            const int i = a - 12;
            const double d = S.D[12 * k + i];
            val += cscale * C.w[i] * d * d + C.sigma;
            if (k < NS - 1) {
              const double e1 = S.E[12 * (k + 1) + i] * d;
              val += R.rho_eq * e1 * e1;
              if (i >= 6) {
                const double e2 = C.dt * S.E[12 * (k + 1) + i - 6] * d;
                val += R.rho_eq * e2 * e2;
              }
            }
          }
        }
        if (a >= 12 && b >= 12 && k < NS - 1) {  // X'(i) X'(i+6) coupling through row (k+1, i)
          const int ia = a - 12, ib2 = b - 12;
          const int lo = ia < ib2 ? ia : ib2, hi = ia < ib2 ? ib2 : ia;
          if (hi - lo == 6) {
            const double e1 = S.E[12 * (k + 1) + lo];
            val += R.rho_eq * C.dt * e1 * e1 * S.D[12 * k + lo] * S.D[12 * k + hi];
          }
        }
        if (a < 12 && b < 12 && a / 3 == b / 3) {  // friction cone rows of this foot
          const int f = a / 3, ca = a - 3 * f, cb = b - 3 * f;
          const int r0 = 384 + 20 * k + 5 * f;
          const double e0 = S.E[r0], e1 = S.E[r0 + 1], e2 = S.E[r0 + 2], e3 = S.E[r0 + 3], e4 = S.E[r0 + 4];
          const double r4 = (25.0 * e4 < 1e-4) ? R.rho_eq : R.rho;
          const double dd = S.D[192 + 12 * k + a] * S.D[192 + 12 * k + b];
          const int lo = ca < cb ? ca : cb, hi = ca < cb ? cb : ca;
          double cg = 0.0;
          if (lo == 0 && hi == 0)
            cg = R.rho * (e0 * e0 + e1 * e1);
          else if (lo == 1 && hi == 1)
            cg = R.rho * (e2 * e2 + e3 * e3);
          else if (lo == 2)
            cg = R.rho * C.mu * C.mu * (e0 * e0 + e1 * e1 + e2 * e2 + e3 * e3) + r4 * e4 * e4;
          else if (lo == 0 && hi == 2)
            cg = R.rho * C.mu * (e1 * e1 - e0 * e0);
          else if (lo == 1 && hi == 2)
            cg = R.rho * C.mu * (e3 * e3 - e2 * e2);
          val += cg * dd;
        }
        if (k > 0) {  // Schur update: - C_k Sinv_{k-1} C_k'
          double sub = 0.0;
          for (int s = 0; s < na; ++s)
            for (int t = 0; t < nb; ++t) sub += ga[s] * N[12 * ra[s] + rb[t]] * gb[t];
          val -= R.rho_eq * R.rho_eq * sub;
        }
        W[e] = val;
      }
    }
    QRW_SYNC();
    // in-place Gauss-Jordan inverse of the SPD block
    for (int p = 0; p < 24; ++p) {
      const double d = 1.0 / W[24 * p + p];
      QRW_LANES(l) {
        if (l < 24) {
          tp[l] = (l == p) ? d : W[24 * p + l] * d;  // scaled pivot row
          tp[24 + l] = W[24 * l + p];                // pivot column
        }
      }
      QRW_SYNC();
      QRW_LANES(l) {
        for (int e = l; e < 576; e += 32) {
          const int r = e / 24, c2 = e - 24 * r;
          double v;
          if (r == p)
            v = tp[c2];
          else if (c2 == p)
            v = -tp[24 + r] * d;
          else
            v = W[e] - tp[24 + r] * tp[c2];
          W[e] = v;
        }
      }
      QRW_SYNC();
    }
    double* Sk = S.Sinv + 300 * k;
    QRW_LANES(l) {
      for (int e = l; e < 300; e += 32) {
        // unpack e -> (a >= b)
        int a = 0;
        while ((a + 1) * (a + 2) / 2 <= e) ++a;
        const int b = e - a * (a + 1) / 2;
        Sk[e] = 0.5 * (W[24 * a + b] + W[24 * b + a]);
      }
    }
    QRW_SYNC();
  }
}

// ------------------------------------------------------------------------------------------
// K x_tilde = rhs.  rhs in wk[0,384), result in wk[384,768).
// ------------------------------------------------------------------------------------------
QRW_HD void mpc_kkt_solve(const MpcConst& C, const MpcSlab& S, const RhoSet R) {
  const double* rhs = S.wk;
  double* xt = S.wk + NX;
  double* tt = S.wk + 768;  // 24
  double* ut = S.wk + 800;  // 12
  // forward: w_k = Sinv_k (rhs_k - C_k w_{k-1})
  for (int k = 0; k < NS; ++k) {
    QRW_LANES(l) {
      if (l < 24) {
        const int a = l;
        double t = rhs[gvar(k, a)];
        if (k > 0) {
          int ra[4];
          const int na = g_rows(a, ra);
          const double* Dp = S.D + 12 * (k - 1);
          const double* wp = xt + 12 * (k - 1);
          double sub = 0.0;
          for (int s = 0; s < na; ++s) {
            const int i = ra[s];
            double jw = Dp[i] * wp[i];
            if (i < 6) jw += C.dt * (Dp[i + 6] * wp[i + 6]);
            sub += g_entry(C, S, k, i, a) * (S.E[12 * k + i] * jw);
          }
          t -= R.rho_eq * sub;
        }
        tt[a] = t;
      }
    }
    QRW_SYNC();
    const double* Sk = S.Sinv + 300 * k;
    QRW_LANES(l) {
      if (l < 24) {
        const int a = l;
        double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
        for (int c2 = 0; c2 < 24; c2 += 4) {
          acc0 += Sk[sym_idx(a, c2)] * tt[c2];
          acc1 += Sk[sym_idx(a, c2 + 1)] * tt[c2 + 1];
          acc2 += Sk[sym_idx(a, c2 + 2)] * tt[c2 + 2];
          acc3 += Sk[sym_idx(a, c2 + 3)] * tt[c2 + 3];
        }
        xt[gvar(k, a)] = (acc0 + acc1) + (acc2 + acc3);
      }
    }
    QRW_SYNC();
  }
  // backward: x_k = w_k - Sinv_k[:, x-part] * (C_{k+1}' x_{k+1})
  for (int k = NS - 2; k >= 0; --k) {
    QRW_LANES(l) {
      if (l < 12) {
        const int p = l;
        // q_i = (scaled dynamics row (k+1, i)) . x_{k+1}  for i = p and (p >= 6) i = p - 6
        double acc = 0.0;
        for (int s = 0; s < ((p >= 6) ? 2 : 1); ++s) {
          const int i = p - 6 * s;
          const int kk = k + 1;
          double q = -(S.D[12 * kk + i] * xt[12 * kk + i]);
          if (i >= 6) {
            const double* Df = S.D + 192 + 12 * kk;
            const double* vf = xt + 192 + 12 * kk;
            if (i < 9) {
              const int c2 = i - 6;
              q += C.dtm * (Df[c2] * vf[c2] + Df[3 + c2] * vf[3 + c2] + Df[6 + c2] * vf[6 + c2] + Df[9 + c2] * vf[9 + c2]);
            } else {
              const double* Br = S.Bang + 36 * kk + 12 * (i - 9);
              double t = 0.0;
#pragma unroll
              for (int a = 0; a < 12; ++a) t += Br[a] * (Df[a] * vf[a]);
              q += t;
            }
          }
          const double e = S.E[12 * kk + i];
          acc += (s ? C.dt : 1.0) * e * (e * q);
        }
        ut[p] = R.rho_eq * S.D[12 * k + p] * acc;
      }
    }
    QRW_SYNC();
    const double* Sk = S.Sinv + 300 * k;
    QRW_LANES(l) {
      if (l < 24) {
        const int a = l;
        double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0;
#pragma unroll
        for (int p = 0; p < 12; p += 3) {
          acc0 += Sk[sym_idx(a, 12 + p)] * ut[p];
          acc1 += Sk[sym_idx(a, 13 + p)] * ut[p + 1];
          acc2 += Sk[sym_idx(a, 14 + p)] * ut[p + 2];
        }
        const int g = gvar(k, a);
        xt[g] = xt[g] - ((acc0 + acc1) + acc2);
      }
    }
    QRW_SYNC();
  }
}

// ------------------------------------------------------------------------------------------
// One full MPC::run for one instance.
//   xref   : 12 x (NS+1) column-major          fsteps : N_gait x 12 column-major
//   st_d/st_i : persistent per-instance state (read unless `first`, always written)
//   out    : 24 x NS column-major x_f_applied  info : {iterations, status, n_factor, flags}
//   dbg    : optional dump (D | E | c | unscaled bounds b) for tests, or nullptr
// ------------------------------------------------------------------------------------------
QRW_HD void mpc_solve_instance(const MpcConst& C, double* slab, const double* xref, const double* fs,
                               double* st_d, int* st_i, int first, double* out, int* info,
                               double* rho_out, double* dbg) {
  const MpcSlab S(slab);
  const int NG = C.N_gait;
  int* gait = S.ib;        // [32] contact bits per gait row
  int* sb = S.ib + 32;     // [16] S_gait bits per step (1 = swing, force pinned to 0)
  int* itmp = S.ib + 48;   // [64] temporaries
  double* red = S.wk + 832;

  // ---- load the state carried between calls
  QRW_LANES(l) {
    for (int i = l; i < NX; i += 32) S.x[i] = first ? 0.0 : st_d[ST_X + i];
    for (int i = l; i < NR; i += 32) {
      S.z[i] = first ? 0.0 : st_d[ST_Z + i];
      S.y[i] = first ? 0.0 : st_d[ST_Y + i];
    }
    for (int i = l; i < 576; i += 32) S.Bang[i] = first ? 0.0 : st_d[ST_BANG + i];
    gait[l] = first ? 0 : st_i[l];
    if (l < NS) sb[l] = first ? 0 : st_i[32 + l];
  }
  double rho_now = first ? C.rho0 : st_d[ST_RHO];

  // ---- construct_gait (MPC.cpp:686-701)
  QRW_LANES(l) {
    int zero = 1, bits = 0;
    if (l < NG) {
      for (int j = 0; j < 12; ++j) {
        const double v = fs[l + NG * j];
        if (!(fabs(v) <= 1e-12)) zero = 0;  // Eigen isZero() with the default 1e-12
      }
      for (int i = 0; i < 4; ++i)
        if (fs[l + NG * (3 * i)] != 0.0) bits |= (1 << i);
    }
    itmp[l] = zero;
    itmp[32 + l] = bits;
  }
  QRW_SYNC();
  int nact = 0;
  while (nact < NG && !itmp[nact]) ++nact;
  const int overrun = (nact >= NG);  // the reference reads/writes out of bounds here
  QRW_LANES(l) {
    if (l < nact)
      gait[l] = itmp[32 + l];
    else if (l == nact && l < NG)
      gait[l] = 0;
  }
  QRW_SYNC();
  // ---- construct_S (MPC.cpp:665-681): rows until the first all-zero gait row
  int nS = 0;
  while (nS < NG && gait[nS] != 0) ++nS;
  QRW_LANES(l) {
    if (l < nS && l < NS) sb[l] = (~gait[l]) & 15;
  }
  // ---- B_k angular blocks (create_ML :213-232 on the first call, update_ML :418-452 after)
  QRW_LANES(l) {
    if (l < NS && (first || l < nS)) {
      const int k = l;
      const double yaw = xref[5 + 12 * k];
      const double cy = cos(yaw), sy = sin(yaw);
      const double Rm[3][3] = {{cy, -sy, 0.0}, {sy, cy, 0.0}, {0.0, 0.0, 1.0}};
      double T[3][3], M[3][3], Ii[3][3];
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
          double a = 0.0;
          for (int q = 0; q < 3; ++q) a += Rm[q][i] * C.gI[3 * q + j];
          T[i][j] = a;
        }
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
          double a = 0.0;
          for (int q = 0; q < 3; ++q) a += T[i][q] * Rm[q][j];
          M[i][j] = a;
        }
      const double c00 = M[1][1] * M[2][2] - M[1][2] * M[2][1];
      const double c10 = M[1][2] * M[2][0] - M[1][0] * M[2][2];
      const double c20 = M[1][0] * M[2][1] - M[1][1] * M[2][0];
      const double det = c00 * M[0][0] + c10 * M[0][1] + c20 * M[0][2];
      const double id = 1.0 / det;
      Ii[0][0] = c00 * id;
      Ii[1][0] = c10 * id;
      Ii[2][0] = c20 * id;
      Ii[0][1] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) * id;
      Ii[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) * id;
      Ii[2][1] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) * id;
      Ii[0][2] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) * id;
      Ii[1][2] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) * id;
      Ii[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) * id;
      for (int i = 0; i < 4; ++i) {
        double v[3];
        for (int r = 0; r < 3; ++r)
          v[r] = first ? (C.foot[4 * r + i] - xref[r + 12 * k])
                       : (fs[k + NG * (3 * i + r)] - (xref[r + 12 * k] + C.off_com[r]));
        const double Sk[3][3] = {{0.0, -v[2], v[1]}, {v[2], 0.0, -v[0]}, {-v[1], v[0], 0.0}};
        for (int r = 0; r < 3; ++r)
          for (int c2 = 0; c2 < 3; ++c2) {
            double a = 0.0;
            for (int q = 0; q < 3; ++q) a += Ii[r][q] * Sk[q][c2];
            S.Bang[36 * k + 12 * r + 3 * i + c2] = C.dt * a;
          }
      }
    }
  }
  // ---- right-hand side of the dynamics rows (create_NK :261-289 / update_NK :469-486)
  QRW_LANES(l) {
    for (int r = l; r < 192; r += 32) {
      const int k = r / 12, i = r - 12 * k;
      double base = (i == 8) ? -C.g8 : 0.0;
      double s = 0.0;
      if (k == 0) {
        double a = 0.0;
        a += 1.0 * (-xref[i]);
        if (i < 6) a += C.dt * (-xref[i + 6]);
        base += a;
      } else {
        s += -1.0 * xref[i + 12 * k];
        if (i < 6) s += -C.dt * xref[i + 6 + 12 * k];
      }
      s += 1.0 * xref[i + 12 * (k + 1)];
      S.bnd[r] = base + s;
    }
  }
  QRW_SYNC();
  if (dbg) {
    QRW_LANES(l) {
      for (int r = l; r < 192; r += 32) dbg[NX + NR + 1 + r] = S.bnd[r];
    }
  }

  // ---- Ruiz equilibration, 10 passes (OSQP scale_data; rerun on every call because
  //      osqp_update_A unscales, overwrites and rescales: MPC.cpp:541,548)
  double cscale = 1.0;
  {
    double* tD = S.Sinv;
    double* tE = S.Sinv + NX;
    QRW_LANES(l) {
      for (int j = l; j < NX; j += 32) S.D[j] = 1.0;
      for (int r = l; r < NR; r += 32) S.E[r] = 1.0;
    }
    QRW_SYNC();
    for (int pass = 0; pass < 10; ++pass) {
      QRW_LANES(l) {
        for (int j = l; j < NX; j += 32) {
          const double d = S.D[j];
          double nrm;
          if (j < 192) {
            const int k = j / 12, i = j - 12 * k;
            nrm = fabs(cscale * C.w[i] * d * d);
            nrm = dmax(nrm, S.E[12 * k + i] * d);
            if (k < NS - 1) {
              nrm = dmax(nrm, S.E[12 * (k + 1) + i] * d);
              if (i >= 6) nrm = dmax(nrm, C.dt * S.E[12 * (k + 1) + i - 6] * d);
            }
          } else {
            const int jj = j - 192, k = jj / 12, a = jj - 12 * k, f = a / 3, c2 = a - 3 * f;
            nrm = cscale * C.wf * d * d;
            nrm = dmax(nrm, C.dtm * S.E[12 * k + 6 + c2] * d);
            for (int r = 0; r < 3; ++r) nrm = dmax(nrm, fabs(S.Bang[36 * k + 12 * r + a]) * S.E[12 * k + 9 + r] * d);
            if ((sb[k] >> f) & 1) nrm = dmax(nrm, S.E[192 + 12 * k + a] * d);
            const int r0 = 384 + 20 * k + 5 * f;
            if (c2 == 0) {
              nrm = dmax(nrm, dmax(S.E[r0], S.E[r0 + 1]) * d);
            } else if (c2 == 1) {
              nrm = dmax(nrm, dmax(S.E[r0 + 2], S.E[r0 + 3]) * d);
            } else {
              const double em = dmax(dmax(S.E[r0], S.E[r0 + 1]), dmax(S.E[r0 + 2], S.E[r0 + 3]));
              nrm = dmax(nrm, C.mu * em * d);
              nrm = dmax(nrm, S.E[r0 + 4] * d);
            }
          }
          tD[j] = 1.0 / sqrt(limit_scaling(nrm));
        }
        for (int r = l; r < NR; r += 32) {
          const double e = S.E[r];
          double nrm;
          if (r < 192) {
            const int k = r / 12, i = r - 12 * k;
            nrm = e * S.D[12 * k + i];
            if (k > 0) {
              nrm = dmax(nrm, e * S.D[12 * (k - 1) + i]);
              if (i < 6) nrm = dmax(nrm, C.dt * e * S.D[12 * (k - 1) + i + 6]);
            }
            if (i >= 6) {
              const double* Df = S.D + 192 + 12 * k;
              if (i < 9) {
                const int c2 = i - 6;
                for (int f = 0; f < 4; ++f) nrm = dmax(nrm, C.dtm * e * Df[3 * f + c2]);
              } else {
                const double* Br = S.Bang + 36 * k + 12 * (i - 9);
                for (int a = 0; a < 12; ++a) nrm = dmax(nrm, fabs(Br[a]) * e * Df[a]);
              }
            }
          } else if (r < 384) {
            const int jj = r - 192, k = jj / 12, a = jj - 12 * k;
            nrm = ((sb[k] >> (a / 3)) & 1) ? e * S.D[r] : 0.0;  // D index of f_k[a] is 192 + 12k + a = r
          } else {
            const int q = r - 384, k = q / 20, rem = q - 20 * k, f = rem / 5, c2 = rem - 5 * f;
            const double* Df = S.D + 192 + 12 * k + 3 * f;
            if (c2 == 4)
              nrm = e * Df[2];
            else
              nrm = dmax(e * Df[c2 >> 1], C.mu * e * Df[2]);
          }
          tE[r] = 1.0 / sqrt(limit_scaling(nrm));
        }
      }
      QRW_SYNC();
      QRW_LANES(l) {
        double part = 0.0;
        for (int j = l; j < NX; j += 32) {
          const double d = S.D[j] * tD[j];
          S.D[j] = d;
          const double p0 = (j < 192) ? C.w[j % 12] : C.wf;
          part += fabs(cscale * p0 * d * d);
        }
        for (int r = l; r < NR; r += 32) S.E[r] = S.E[r] * tE[r];
        red[l] = part;
      }
      QRW_SYNC();
      double c_temp = red_sum32(red) / (double)NX;
      c_temp = dmax(c_temp, 1.0);  // ||q||_inf = 0 -> limit_scaling -> 1
      c_temp = limit_scaling(c_temp);
      cscale *= 1.0 / c_temp;
      QRW_SYNC();
    }
  }
  const double cinv = 1.0 / cscale;
  if (dbg) {
    QRW_LANES(l) {
      for (int j = l; j < NX; j += 32) dbg[j] = S.D[j];
      for (int r = l; r < NR; r += 32) dbg[NX + r] = S.E[r];
      if (l == 0) dbg[NX + NR] = cscale;
    }
  }
  QRW_LANES(l) {
    for (int r = l; r < 192; r += 32) S.bnd[r] = S.E[r] * S.bnd[r];
  }
  QRW_SYNC();

  // ---- KKT factorisation
  RhoSet R(rho_now);
  int n_factor = 1;
  mpc_factor(C, S, R, cscale);

  // ---- ADMM (osqp_solve)
  double* rhs = S.wk;
  double* xt = S.wk + NX;
  int status = ST_UNSOLVED;
  int iter = 0;
  for (iter = 1; iter <= C.max_iter; ++iter) {
    // rhs = sigma x + A'(rho z - y)          (q = 0 for the MPC, MPC.cpp:385)
    QRW_LANES(l) {
      auto u = [&](int r) {
        const double rr = row_is_eq(S, r) ? R.rho_eq : R.rho;
        return S.E[r] * (rr * S.z[r] - S.y[r]);
      };
      for (int j = l; j < 192; j += 32) {
        const int k = j / 12, i = j - 12 * k;
        rhs[j] = C.sigma * S.x[j] + S.D[j] * a0t_col_state(C, k, i, u);
      }
      for (int j = 192 + l; j < NX; j += 32) {
        const int jj = j - 192, k = jj / 12, a = jj - 12 * k;
        rhs[j] = C.sigma * S.x[j] + S.D[j] * a0t_col_force(C, S.Bang, sb, k, a, u);
      }
    }
    QRW_SYNC();
    mpc_kkt_solve(C, S, R);
    // z_tilde = A x_tilde, relaxation, projection, dual update
    QRW_LANES(l) {
      for (int r = l; r < 192; r += 32) {  // dynamics rows: l = u = bnd
        const int k = r / 12, i = r - 12 * k;
        const double zt = S.E[r] * a0_row_dyn(C, S.D, S.Bang, xt, k, i);
        const double zr = C.alpha * zt + (1.0 - C.alpha) * S.z[r];
        const double zn = S.bnd[r];
        S.y[r] += R.rho_eq * (zr - zn);
        S.z[r] = zn;
      }
      for (int r = 192 + l; r < 384; r += 32) {  // force-enable rows: l = u = 0
        const int jj = r - 192, k = jj / 12, a = jj - 12 * k;
        const double zt = ((sb[k] >> (a / 3)) & 1) ? S.E[r] * (S.D[r] * xt[r]) : 0.0;
        const double zr = C.alpha * zt + (1.0 - C.alpha) * S.z[r];
        S.y[r] += R.rho_eq * (zr - 0.0);
        S.z[r] = 0.0;
      }
      for (int r = 384 + l; r < NR; r += 32) {  // cone rows: [-inf | -25 E, 0]
        const int q = r - 384;
        const bool eq = row_is_eq(S, r);
        const double rr = eq ? R.rho_eq : R.rho, ri = eq ? R.rinv_eq : R.rinv;
        const double zt = S.E[r] * a0_row_cone(C, S.D, xt, q);
        const double zr = C.alpha * zt + (1.0 - C.alpha) * S.z[r];
        double zn = zr + ri * S.y[r];
        if (q % 5 == 4) zn = dmax(zn, -25.0 * S.E[r]);
        zn = dmin(zn, 0.0);
        S.y[r] += rr * (zr - zn);
        S.z[r] = zn;
      }
      for (int j = l; j < NX; j += 32) S.x[j] = C.alpha * xt[j] + (1.0 - C.alpha) * S.x[j];
    }
    QRW_SYNC();

    if (iter % C.check_every == 0) {
      // residuals (auxil.c compute_pri_res / compute_dua_res / compute_*_tol)
      double* rp = S.wk;  // 12 x 32 partial maxima
      QRW_LANES(l) {
        double m0 = 0, m1 = 0, m2 = 0, m3 = 0, m4 = 0, m5 = 0;
        for (int r = l; r < NR; r += 32) {
          double ax;
          if (r < 192) {
            const int k = r / 12;
            ax = S.E[r] * a0_row_dyn(C, S.D, S.Bang, S.x, k, r - 12 * k);
          } else if (r < 384) {
            const int jj = r - 192, k = jj / 12, a = jj - 12 * k;
            ax = ((sb[k] >> (a / 3)) & 1) ? S.E[r] * (S.D[r] * S.x[r]) : 0.0;
          } else {
            ax = S.E[r] * a0_row_cone(C, S.D, S.x, r - 384);
          }
          const double zi = S.z[r], einv = 1.0 / S.E[r], rs = ax - zi;
          m0 = dmax(m0, fabs(rs));
          m1 = dmax(m1, fabs(einv * rs));
          m2 = dmax(m2, fabs(zi));
          m3 = dmax(m3, fabs(einv * zi));
          m4 = dmax(m4, fabs(ax));
          m5 = dmax(m5, fabs(einv * ax));
        }
        double n0 = 0, n1 = 0, n2 = 0, n3 = 0, n4 = 0, n5 = 0;
        auto uy = [&](int r) { return S.E[r] * S.y[r]; };
        for (int j = l; j < NX; j += 32) {
          double aty, p0;
          if (j < 192) {
            const int k = j / 12, i = j - 12 * k;
            aty = S.D[j] * a0t_col_state(C, k, i, uy);
            p0 = C.w[i];
          } else {
            const int jj = j - 192, k = jj / 12, a = jj - 12 * k;
            aty = S.D[j] * a0t_col_force(C, S.Bang, sb, k, a, uy);
            p0 = C.wf;
          }
          const double px = (cscale * p0 * S.D[j] * S.D[j]) * S.x[j];
          const double ds = px + aty, dinv = 1.0 / S.D[j];
          n0 = dmax(n0, fabs(ds));
          n1 = dmax(n1, fabs(dinv * ds));
          n2 = dmax(n2, fabs(aty));
          n3 = dmax(n3, fabs(dinv * aty));
          n4 = dmax(n4, fabs(px));
          n5 = dmax(n5, fabs(dinv * px));
        }
        rp[l] = m0;
        rp[32 + l] = m1;
        rp[64 + l] = m2;
        rp[96 + l] = m3;
        rp[128 + l] = m4;
        rp[160 + l] = m5;
        rp[192 + l] = n0;
        rp[224 + l] = n1;
        rp[256 + l] = n2;
        rp[288 + l] = n3;
        rp[320 + l] = n4;
        rp[352 + l] = n5;
      }
      QRW_SYNC();
      const double pri_s = red_max32(rp), pri_u = red_max32(rp + 32), nz_s = red_max32(rp + 64),
                   nz_u = red_max32(rp + 96), nax_s = red_max32(rp + 128), nax_u = red_max32(rp + 160),
                   dua_s = red_max32(rp + 192), dua_u = cinv * red_max32(rp + 224),
                   naty_s = red_max32(rp + 256), naty_u = red_max32(rp + 288),
                   npx_s = red_max32(rp + 320), npx_u = red_max32(rp + 352);
      QRW_SYNC();
      if (pri_u > 1e30 || dua_u > 1e30) {
        status = ST_NON_CVX;
        break;
      }
      const double eps_pri = C.eps_abs + C.eps_rel * dmax(nz_u, nax_u);
      const double eps_dua = C.eps_abs + C.eps_rel * (cinv * dmax(naty_u, npx_u));
      // Primal / dual infeasibility can never be reported for this QP: with l = -inf on cone rows
      // OSQP's certificate sum is NaN (is_primal_infeasible), and q = 0 makes q'dx < 0 false.
      if (pri_u < eps_pri && dua_u < eps_dua) {
        status = ST_SOLVED;
        break;
      }
      if (iter == C.max_iter) {  // check_termination(work, 1): tolerances x 10
        const double e1 = 10.0 * C.eps_abs + 10.0 * C.eps_rel * dmax(nz_u, nax_u);
        const double e2 = 10.0 * C.eps_abs + 10.0 * C.eps_rel * (cinv * dmax(naty_u, npx_u));
        status = (pri_u < e1 && dua_u < e2) ? ST_SOLVED_INACCURATE : ST_MAX_ITER;
      }
      if (iter % C.adapt_every == 0) {  // adapt_rho / compute_rho_estimate (scaled-space norms)
        const double pr = pri_s / (dmax(nz_s, nax_s) + 1e-10);
        const double dr = dua_s / (dmax(naty_s, npx_s) + 1e-10);
        double est = R.rho * sqrt(pr / (dr + 1e-10));
        est = dmin(dmax(est, 1e-6), 1e6);
        if (est > R.rho * C.adapt_tol || est < R.rho / C.adapt_tol) {
          R = RhoSet(est);
          mpc_factor(C, S, R, cscale);
          ++n_factor;
        }
      }
    }
  }
  if (iter > C.max_iter) iter = C.max_iter;
  if (status == ST_UNSOLVED) status = ST_MAX_ITER;

  // ---- retrieve_result (MPC.cpp:569-599) + state carried to the next call
  const bool bad = (status == ST_NON_CVX);
  QRW_LANES(l) {
    for (int e = l; e < 12 * NS; e += 32) {
      const int i = e / 12, k = e - 12 * i;
      const double nanv = nan("");
      out[k + 24 * i] = bad ? nanv : (S.D[12 * i + k] * S.x[12 * i + k] + xref[k + 12 * (1 + i)]);
      out[12 + k + 24 * i] = bad ? nanv : (S.D[192 + 12 * i + k] * S.x[192 + 12 * i + k]);
    }
    for (int i = l; i < NX; i += 32) st_d[ST_X + i] = bad ? 0.0 : S.x[i];
    for (int i = l; i < NR; i += 32) {
      st_d[ST_Z + i] = bad ? 0.0 : S.z[i];
      st_d[ST_Y + i] = bad ? 0.0 : S.y[i];
    }
    for (int i = l; i < 576; i += 32) st_d[ST_BANG + i] = S.Bang[i];
    st_i[l] = gait[l];
    if (l < NS) st_i[32 + l] = sb[l];
    if (l == 0) {
      st_d[ST_RHO] = R.rho;
      info[0] = iter;
      info[1] = status;
      info[2] = n_factor;
      info[3] = overrun;
      if (rho_out) *rho_out = R.rho;
    }
  }
  QRW_SYNC();
}

}  // namespace qrw
