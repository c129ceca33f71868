// mpc_cta_core.cuh -- one Solo12 centroidal-MPC solve executed by ONE 256-thread CTA.
//
// Same contract, same arithmetic model and same persistent state as mpc_core.cuh (the one-warp
// version this kernel supersedes on the throughput path); it replaces the same reference code:
//   MPC::run  /root/reference/src/MPC.cpp:626-649  (construct_gait :686, construct_S :665, update_ML :418,
//   update_NK :469, call_solver :501 -> OSQP, retrieve_result :569).
//
// Why a CTA per instance.  ncu on the one-warp kernel (profiles/r01_ncu_mpc_v0_*.txt) shows it is
// bound by the ~12.6 k instructions ONE warp has to issue per ADMM iteration (issue slot 22 % busy, 4.6
// cycles per instruction, FP64 pipe 6 %).  Here the work of an iteration is spread thread-per-variable:
//   * warp w owns the horizon stages k = 2w, 2w+1; lane l < 24 owns stage variable l of both stages
//     (l < 12: force f_k[l], l >= 12: state X_{k+1}[l-12]); lane l < 12 also owns the dynamics row
//     (k, l) and the force-enable row (k, l), lane l >= 12 the friction-cone row (k, l-12).  Iterates
//     x / z / y and the hot scaling entries live in REGISTERS of their owner.
//   * the block-tridiagonal KKT solve (24x24 pivot-block inverses S_k) is re-associated so that the only
//     sequential part is a 12-vector recurrence:
//         forward   p_{k+1} = p^_{k+1} - T_k' p_k ,   p^_{k+1} = Jx_{k+1} [S_k t_k]_X
//         backward  g_k     = g^_k     - T_k  g_{k+1}, g^_k    = G_k S_k (t_k - rho G_k' p_k)
//     with T_k = rho G_k S_k[:,X] Jx_{k+1}' (12 x 12, built at factorisation time).  Everything else
//     (three applications of S_k per stage, the constraint-matrix products, projection, dual update)
//     runs on all 16 stages at once.  G_k / Jx_k are the scaled dynamics rows of stage k restricted to the
//     stage's own variables / to X_k.
//   * S_k is split: its 12 state columns sit in registers of the row owner (24 doubles per thread for
//     two stages), its 12 force columns in shared memory, column-major (conflict free); that halves
//     the shared-memory traffic that bounds an all-in-shared layout (DESIGN.md "MPC kernel").
// Shared memory: 84 KB per CTA -> two CTAs (16 warps) per SM, <= 128 registers per thread.
#pragma once
#include "mpc_core.cuh"

namespace qrw {
namespace cta {

constexpr int NT = 256;   // threads per CTA
constexpr int SPW = 2;    // stages per warp

// shared slab layout (doubles)
constexpr int C_SF = 0;                    // [16][12][24]  S_k(:, force columns), column-major
constexpr int C_TT = C_SF + NS * 288;      // [14][12][13]  T_k, k = 1..14, rows padded to 13
constexpr int C_BANG = C_TT + 14 * 156;    // [16][3][12]   angular blocks of B_k (unscaled)
constexpr int C_D = C_BANG + 576;          // [16][24]      Ruiz column scaling, stage-local order
constexpr int C_E = C_D + 384;             // [16][44]      Ruiz row scaling: 12 dyn | 12 enable | 20 cone
constexpr int C_U = C_E + 704;             // [16][44]      row-space exchange
constexpr int C_TV = C_U + 704;            // [16][24]      right-hand side of the stage solves
constexpr int C_DX = C_TV + 384;           // [16][24]      D .* vector, variable-space exchange
constexpr int C_PH = C_DX + 384;           // [16][12]      p^ -> p
constexpr int C_GH = C_PH + 192;           // [16][12]      g^ -> g
constexpr int C_UX = C_GH + 192;           // [16][12]
constexpr int C_RED = C_UX + 192;          // 128           reductions / broadcast scalars
constexpr int C_INT = C_RED + 128;         // 128 ints
constexpr int CTA_SLAB_DOUBLES = C_INT + 64;
constexpr int CTA_SLAB_BYTES = CTA_SLAB_DOUBLES * 8;
// factorisation workspace, aliased on U | TV | DX (1472 doubles): W0, W1 (24x24 ping-pong), N (12x12)
constexpr int C_W0 = C_U, C_W1 = C_U + 576, C_N = C_U + 1152;
// Ruiz cost-normalisation terms, aliased on U (384 doubles, global variable order)
constexpr int C_TMP = C_U;

struct Sm {
  double *SF, *TT, *Bang, *D, *E, *U, *TV, *DX, *PH, *GH, *UX, *red;
  int* ib;
  QRW_HD explicit Sm(double* b)
      : SF(b + C_SF), TT(b + C_TT), Bang(b + C_BANG), D(b + C_D), E(b + C_E), U(b + C_U), TV(b + C_TV),
        DX(b + C_DX), PH(b + C_PH), GH(b + C_GH), UX(b + C_UX), red(b + C_RED),
        ib(reinterpret_cast<int*>(b + C_INT)) {}
};

// global row index of stage-local row r of stage k (12 dyn | 12 enable | 20 cone)
QRW_HD int grow(int k, int r) { return r < 12 ? 12 * k + r : (r < 24 ? 192 + 12 * k + (r - 12) : 384 + 20 * k + (r - 24)); }

// (unscaled dynamics rows of stage k restricted to the stage's force variable a)' * v, v = 12 row values
QRW_HD double dynT_force(const MpcConst& C, const double* Bk, const double* v, int a, int c3) {
  double s = C.dtm * v[6 + c3];
  s += Bk[a] * v[9];
  s += Bk[12 + a] * v[10];
  s += Bk[24 + a] * v[11];
  return s;
}
// unscaled dynamics row i of stage k applied to (D .* vector) of the stage's own variables (dx, 24)
QRW_HD double dyn_row_own(const MpcConst& C, const double* Bk, const double* dx, int i) {
  double s = -dx[12 + i];
  if (i >= 6) {
    if (i < 9) {
      const int c = i - 6;
      s += C.dtm * (dx[c] + dx[3 + c] + dx[6 + c] + dx[9 + c]);
    } else {
      const double* Br = Bk + 12 * (i - 9);
      double t = 0.0;
#pragma unroll
      for (int a = 0; a < 12; ++a) t += Br[a] * dx[a];
      s += t;
    }
  }
  return s;
}
// unscaled friction-cone row q of a stage applied to (D .* vector) of the stage's force variables
QRW_HD double cone_row(const MpcConst& C, const double* dx, int cf, int cc) {
  const double fz = dx[3 * cf + 2];
  if (cc == 4) return -fz;
  const double ft = dx[3 * cf + (cc >> 1)];
  return ((cc & 1) ? -ft : ft) - C.mu * fz;
}

// scaled dynamics-row entry of row (k,i) on stage variable a (same as g_entry of mpc_core.cuh)
QRW_HD double gent(const MpcConst& C, const Sm& S, int k, int i, int a) {
  const double e = S.E[44 * k + i];
  if (a >= 12) return (a - 12 == i) ? -(e * S.D[24 * k + a]) : 0.0;
  if (i < 6) return 0.0;
  const double ed = e * S.D[24 * k + a];
  if (i < 9) return (a % 3 == i - 6) ? C.dtm * ed : 0.0;
  return S.Bang[36 * k + 12 * (i - 9) + a] * ed;
}

#if defined(QRW_DEBUG_SOLVE) && !QRW_ON_GPU
inline double g_dbgK[16][576];
#endif
struct Regs {
  double x[SPW], z1[SPW], y1[SPW], y2[SPW], e1[SPW], b1[SPW], d[SPW];
  double sx[SPW][12];  // S_k(l, state columns)
};

// ------------------------------------------------------------------------------------------
// Block LDL' of K = P + sigma I + A' diag(rho) A : pivot-block inverses S_k -> SF (shared) + R.sx
// (registers), recurrence matrices T_k -> TT.  All threads of the CTA.
// ------------------------------------------------------------------------------------------
QRW_HD void factor(const MpcConst& C, const Sm& S, Regs& R, const RhoSet Rh, double cscale, int tid) {
  double* W0 = S.U;  // aliases, see C_W0 ...
  double* W1 = S.U + 576;
  double* N = S.U + 1152;
  const int* sb = S.ib + 32;
  const int w = tid >> 5, l = tid & 31;
  for (int k = 0; k < NS; ++k) {
    // ---- assemble the 24 x 24 diagonal block of stage k minus the Schur term (N from stage k-1)
    for (int e = tid; e < 576; e += NT) {
      const int a = e / 24, b = e - 24 * a;
      int ra[4], rb[4];
      const int na = g_rows(a, ra), nb = g_rows(b, rb);
      double ga[4], gb[4];
      for (int s = 0; s < na; ++s) ga[s] = gent(C, S, k, ra[s], a);
      for (int t = 0; t < nb; ++t) gb[t] = gent(C, S, k, rb[t], b);
      double val = 0.0;
      for (int s = 0; s < na; ++s)
        for (int t = 0; t < nb; ++t)
          if (ra[s] == rb[t]) val += ga[s] * gb[t];
      val *= Rh.rho_eq;
      if (a == b) {
        const double d = S.D[24 * k + a];
        if (a < 12) {
          val += cscale * C.wf * d * d + C.sigma;
          if ((sb[k] >> (a / 3)) & 1) {
            const double sd = S.E[44 * k + 12 + a] * d;
            val += Rh.rho_eq * sd * sd;
          }
        } else {
          const int i = a - 12;
          val += cscale * C.w[i] * d * d + C.sigma;
          if (k < NS - 1) {
            const double e1 = S.E[44 * (k + 1) + i] * d;
            val += Rh.rho_eq * e1 * e1;
            if (i >= 6) {
              const double e2 = C.dt * S.E[44 * (k + 1) + i - 6] * d;
              val += Rh.rho_eq * e2 * e2;
            }
          }
        }
      }
      if (a >= 12 && b >= 12 && k < NS - 1) {  // X(i) X(i+6) coupling through row (k+1, i)
        const int ia = a - 12, ib2 = b - 12;
        const int lo = ia < ib2 ? ia : ib2, hi = ia < ib2 ? ib2 : ia;
        if (hi - lo == 6) {
          const double e1 = S.E[44 * (k + 1) + lo];
          val += Rh.rho_eq * C.dt * e1 * e1 * S.D[24 * k + 12 + lo] * S.D[24 * k + 12 + hi];
        }
      }
      if (a < 12 && b < 12 && a / 3 == b / 3) {  // friction-cone rows of this foot
        const int f = a / 3, ca = a - 3 * f, cb = b - 3 * f;
        const double* Ec = S.E + 44 * k + 24 + 5 * f;
        const double e0 = Ec[0], e1 = Ec[1], e2 = Ec[2], e3 = Ec[3], e4 = Ec[4];
        const double r4 = (25.0 * e4 < 1e-4) ? Rh.rho_eq : Rh.rho;
        const double dd = S.D[24 * k + a] * S.D[24 * k + b];
        const int lo = ca < cb ? ca : cb, hi = ca < cb ? cb : ca;
        double cg = 0.0;
        if (lo == 0 && hi == 0)
          cg = Rh.rho * (e0 * e0 + e1 * e1);
        else if (lo == 1 && hi == 1)
          cg = Rh.rho * (e2 * e2 + e3 * e3);
        else if (lo == 2)
          cg = Rh.rho * C.mu * C.mu * (e0 * e0 + e1 * e1 + e2 * e2 + e3 * e3) + r4 * e4 * e4;
        else if (lo == 0 && hi == 2)
          cg = Rh.rho * C.mu * (e1 * e1 - e0 * e0);
        else if (lo == 1 && hi == 2)
          cg = Rh.rho * C.mu * (e3 * e3 - e2 * e2);
        val += cg * dd;
      }
      if (k > 0) {
        double sub = 0.0;
        for (int s = 0; s < na; ++s)
          for (int t = 0; t < nb; ++t) sub += ga[s] * N[12 * ra[s] + rb[t]] * gb[t];
        val -= Rh.rho_eq * Rh.rho_eq * sub;
      }
      W0[e] = val;
#if defined(QRW_DEBUG_SOLVE) && !QRW_ON_GPU
      g_dbgK[k][e] = val;
#endif
    }
    bar();
    // ---- Gauss-Jordan inverse of the SPD block, ping-pong between W0 and W1 (one barrier per pivot)
    for (int p = 0; p < 24; ++p) {
      const double* Wc = (p & 1) ? W1 : W0;
      double* Wn = (p & 1) ? W0 : W1;
      const double d = 1.0 / Wc[24 * p + p];
      for (int e = tid; e < 576; e += NT) {
        const int r = e / 24, c2 = e - 24 * r;
        double v;
        if (r == p)
          v = (c2 == p) ? d : Wc[24 * p + c2] * d;
        else if (c2 == p)
          v = -Wc[24 * r + p] * d;
        else
          v = Wc[e] - Wc[24 * r + p] * (Wc[24 * p + c2] * d);
        Wn[e] = v;
      }
      bar();
    }
    // ---- distribute the (symmetrised) inverse: force columns -> shared, state columns -> registers
    for (int e = tid; e < 288; e += NT) {
      const int c = e / 24, a = e - 24 * c;
      S.SF[288 * k + e] = 0.5 * (W0[24 * a + c] + W0[24 * c + a]);
    }
    if (l < 24) {
#pragma unroll
      for (int s = 0; s < SPW; ++s)
        if (k == SPW * w + s) {
#pragma unroll
          for (int j = 0; j < 12; ++j) R.sx[s][j] = 0.5 * (W0[24 * l + 12 + j] + W0[24 * (12 + j) + l]);
        }
    }
    // ---- T_k = rho G_k S_k(:,X) Jx_{k+1}'   (needed for k = 1..14)
    if (k >= 1 && k <= NS - 2 && tid < 144) {
      const int i = tid / 12, j = tid - 12 * i;
      const double ej = S.E[44 * (k + 1) + j];
      const double dj = S.D[24 * k + 12 + j];
      const double dj6 = (j < 6) ? C.dt * S.D[24 * k + 18 + j] : 0.0;
      // V(a) = E_{k+1,j} (S(a,12+j) D_j + dt S(a,18+j) D_{j+6})
      auto V = [&](int a) {
        double v = 0.5 * (W0[24 * a + 12 + j] + W0[24 * (12 + j) + a]) * dj;
        if (j < 6) v += 0.5 * (W0[24 * a + 18 + j] + W0[24 * (18 + j) + a]) * dj6;
        return ej * v;
      };
      double acc = gent(C, S, k, i, 12 + i) * V(12 + i);
      if (i >= 6) {
        if (i < 9) {
          for (int f = 0; f < 4; ++f) acc += gent(C, S, k, i, 3 * f + (i - 6)) * V(3 * f + (i - 6));
        } else {
          for (int a = 0; a < 12; ++a) acc += gent(C, S, k, i, a) * V(a);
        }
      }
      S.TT[156 * (k - 1) + 13 * i + j] = Rh.rho_eq * acc;
    }
    // ---- N = Jx_{k+1} S_k(X,X) Jx_{k+1}' for the next stage's Schur term
    if (k < NS - 1 && tid >= 160 && tid < 160 + 72) {
      // two entries per thread
      for (int e = (tid - 160) * 2; e < (tid - 160) * 2 + 2; ++e) {
        const int i = e / 12, j = e - 12 * i;
        const double ei = S.E[44 * (k + 1) + i], ej = S.E[44 * (k + 1) + j];
        const double* Dp = S.D + 24 * k + 12;
        double acc = 0.0;
        const int ni = (i < 6) ? 2 : 1, nj = (j < 6) ? 2 : 1;
        for (int s = 0; s < ni; ++s) {
          const int p = i + 6 * s;
          const double cp = (s ? C.dt : 1.0) * ei * Dp[p];
          for (int t = 0; t < nj; ++t) {
            const int q = j + 6 * t;
            const double cq = (t ? C.dt : 1.0) * ej * Dp[q];
            acc += cp * (0.5 * (W0[24 * (12 + p) + 12 + q] + W0[24 * (12 + q) + 12 + p])) * cq;
          }
        }
        N[e] = acc;
      }
    }
    bar();
  }
}

// ------------------------------------------------------------------------------------------
// One full MPC::run for one instance; every thread of the CTA calls it with the same arguments.
// Same I/O contract as mpc_solve_instance (mpc_core.cuh).
// ------------------------------------------------------------------------------------------
QRW_HD void mpc_cta_solve(const MpcConst& C, double* slab, const double* xref, const double* fs, double* st_d,
                          int* st_i, int first, double* out, int* info, double* rho_out, double* dbg) {
  const Sm S(slab);
  const int tid = cta::tid();
  const int w = tid >> 5, l = tid & 31;
  const int NG = C.N_gait;
  int* gait = S.ib;       // [32]
  int* sb = S.ib + 32;    // [16] S_gait bits per step (bit f set = foot f swings, force pinned to 0)
  int* itmp = S.ib + 48;  // [64]
  int* ictl = S.ib + 112; // [16] broadcast control words
  Regs R;
  // lane roles
  const bool isvar = l < 24, isF = l < 12, isrow2 = l < 12;
  const int a = l;                      // stage variable
  const int af = isF ? a / 3 : 0, ac = isF ? a - 3 * af : 0;
  const int xi = l - 12;                // state component for 12 <= l < 24
  const int cq = l - 12;                // cone row for l >= 12
  const int cf = (l >= 12) ? cq / 5 : 0, cc = (l >= 12) ? cq - 5 * cf : 0;
  const int r1 = isrow2 ? l : 12 + l;   // stage-local index of the first owned row (dyn | cone)

  // ---- load the state carried between calls
  for (int i = tid; i < 576; i += NT) S.Bang[i] = first ? 0.0 : st_d[ST_BANG + i];
  if (tid < 32) gait[tid] = first ? 0 : st_i[tid];
  if (tid < NS) sb[tid] = first ? 0 : st_i[32 + tid];
  if (tid < 12) S.PH[tid] = 0.0;  // p_0 = 0
#pragma unroll
  for (int s = 0; s < SPW; ++s) {
    const int k = SPW * w + s;
    R.x[s] = (isvar && !first) ? st_d[ST_X + gvar(k, a)] : 0.0;
    const int g1 = grow(k, r1);
    R.z1[s] = first ? 0.0 : st_d[ST_Z + g1];
    R.y1[s] = first ? 0.0 : st_d[ST_Y + g1];
    R.y2[s] = (isrow2 && !first) ? st_d[ST_Y + grow(k, 12 + l)] : 0.0;
  }
  double rho_now = first ? C.rho0 : st_d[ST_RHO];

  // ---- construct_gait (MPC.cpp:686-701), construct_S (:665-681): warp 0, as in mpc_core.cuh
  if (w == 0) {
    int zero = 1, bits = 0;
    if (l < NG) {
      for (int j = 0; j < 12; ++j) {
        const double v = fs[l + NG * j];
        if (!(fabs(v) <= 1e-12)) zero = 0;
      }
      for (int i = 0; i < 4; ++i)
        if (fs[l + NG * (3 * i)] != 0.0) bits |= (1 << i);
    }
    itmp[l] = zero;
    itmp[32 + l] = bits;
  }
  bar();
  int nact = 0;
  while (nact < NG && !itmp[nact]) ++nact;
  const int overrun = (nact >= NG);
  bar();
  if (w == 0) {
    if (l < nact)
      gait[l] = itmp[32 + l];
    else if (l == nact && l < NG)
      gait[l] = 0;
  }
  bar();
  int nS = 0;
  while (nS < NG && gait[nS] != 0) ++nS;
  bar();
  if (tid < nS && tid < NS) sb[tid] = (~gait[tid]) & 15;
  // ---- B_k angular blocks (create_ML :213-232 on the first call, update_ML :418-452 after)
  if (tid < NS && (first || tid < nS)) {
    const int k = tid;
    const double yaw = xref[5 + 12 * k];
    const double cy = cos(yaw), sy = sin(yaw);
    const double Rm[3][3] = {{cy, -sy, 0.0}, {sy, cy, 0.0}, {0.0, 0.0, 1.0}};
    double T[3][3], M[3][3], Ii[3][3];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        double acc = 0.0;
        for (int q = 0; q < 3; ++q) acc += Rm[q][i] * C.gI[3 * q + j];
        T[i][j] = acc;
      }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        double acc = 0.0;
        for (int q = 0; q < 3; ++q) acc += T[i][q] * Rm[q][j];
        M[i][j] = acc;
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
          double acc = 0.0;
          for (int q = 0; q < 3; ++q) acc += Ii[r][q] * Sk[q][c2];
          S.Bang[36 * k + 12 * r + 3 * i + c2] = C.dt * acc;
        }
    }
  }
  // ---- right-hand side of the dynamics rows (create_NK :261-289 / update_NK :469-486), row owners
#pragma unroll
  for (int s = 0; s < SPW; ++s) {
    const int k = SPW * w + s;
    double bnd = 0.0;
    if (isrow2) {
      const int i = l;
      double base = (i == 8) ? -C.g8 : 0.0;
      double sm = 0.0;
      if (k == 0) {
        double acc = 0.0;
        acc += 1.0 * (-xref[i]);
        if (i < 6) acc += C.dt * (-xref[i + 6]);
        base += acc;
      } else {
        sm += -1.0 * xref[i + 12 * k];
        if (i < 6) sm += -C.dt * xref[i + 6 + 12 * k];
      }
      sm += 1.0 * xref[i + 12 * (k + 1)];
      bnd = base + sm;
      if (dbg) dbg[NX + NR + 1 + 12 * k + i] = bnd;
    }
    R.b1[s] = bnd;
    // Ruiz start values
    R.d[s] = 1.0;
    R.e1[s] = 1.0;
    if (isvar) S.D[24 * k + a] = 1.0;
    S.E[44 * k + r1] = 1.0;
    if (isrow2) S.E[44 * k + 12 + l] = 1.0;
  }
  bar();

  // ---- Ruiz equilibration, 10 passes (OSQP scale_data; rerun on every call, MPC.cpp:541,548)
  double cscale = 1.0;
  {
    double e2[SPW];
#pragma unroll
    for (int s = 0; s < SPW; ++s) e2[s] = 1.0;
    for (int pass = 0; pass < 10; ++pass) {
      double dD[SPW], dE1[SPW], dE2[SPW];
#pragma unroll
      for (int s = 0; s < SPW; ++s) {
        const int k = SPW * w + s;
        const double* Ek = S.E + 44 * k;
        const double* Dk = S.D + 24 * k;
        dD[s] = dE1[s] = dE2[s] = 1.0;
        if (isvar) {
          const double d = R.d[s];
          double nrm;
          if (!isF) {
            const int i = xi;
            nrm = fabs(cscale * C.w[i] * d * d);
            nrm = dmax(nrm, Ek[i] * d);
            if (k < NS - 1) {
              nrm = dmax(nrm, Ek[44 + i] * d);
              if (i >= 6) nrm = dmax(nrm, C.dt * Ek[44 + i - 6] * d);
            }
          } else {
            nrm = cscale * C.wf * d * d;
            nrm = dmax(nrm, C.dtm * Ek[6 + ac] * d);
            for (int r = 0; r < 3; ++r) nrm = dmax(nrm, fabs(S.Bang[36 * k + 12 * r + a]) * Ek[9 + r] * d);
            if ((sb[k] >> af) & 1) nrm = dmax(nrm, Ek[12 + a] * d);
            const double* Ec = Ek + 24 + 5 * af;
            if (ac == 0) {
              nrm = dmax(nrm, dmax(Ec[0], Ec[1]) * d);
            } else if (ac == 1) {
              nrm = dmax(nrm, dmax(Ec[2], Ec[3]) * d);
            } else {
              const double em = dmax(dmax(Ec[0], Ec[1]), dmax(Ec[2], Ec[3]));
              nrm = dmax(nrm, C.mu * em * d);
              nrm = dmax(nrm, Ec[4] * d);
            }
          }
          dD[s] = 1.0 / sqrt(limit_scaling(nrm));
        }
        if (isrow2) {
          const int i = l;
          const double e = R.e1[s];
          double nrm = e * Dk[12 + i];
          if (k > 0) {
            nrm = dmax(nrm, e * Dk[-24 + 12 + i]);
            if (i < 6) nrm = dmax(nrm, C.dt * e * Dk[-24 + 18 + i]);
          }
          if (i >= 6) {
            if (i < 9) {
              for (int f = 0; f < 4; ++f) nrm = dmax(nrm, C.dtm * e * Dk[3 * f + i - 6]);
            } else {
              const double* Br = S.Bang + 36 * k + 12 * (i - 9);
              for (int b = 0; b < 12; ++b) nrm = dmax(nrm, fabs(Br[b]) * e * Dk[b]);
            }
          }
          dE1[s] = 1.0 / sqrt(limit_scaling(nrm));
          const double nf = ((sb[k] >> af) & 1) ? e2[s] * Dk[a] : 0.0;
          dE2[s] = 1.0 / sqrt(limit_scaling(nf));
        } else {
          const double e = R.e1[s];
          const double* Df = Dk + 3 * cf;
          double nrm;
          if (cc == 4)
            nrm = e * Df[2];
          else
            nrm = dmax(e * Df[cc >> 1], C.mu * e * Df[2]);
          dE1[s] = 1.0 / sqrt(limit_scaling(nrm));
        }
      }
      bar();
#pragma unroll
      for (int s = 0; s < SPW; ++s) {
        const int k = SPW * w + s;
        if (isvar) {
          const double d = R.d[s] * dD[s];
          R.d[s] = d;
          S.D[24 * k + a] = d;
          const double p0 = isF ? C.wf : C.w[xi];
          S.U[gvar(k, a)] = fabs(cscale * p0 * d * d);  // C_TMP, global variable order
        }
        R.e1[s] *= dE1[s];
        S.E[44 * k + r1] = R.e1[s];
        if (isrow2) {
          e2[s] *= dE2[s];
          S.E[44 * k + 12 + l] = e2[s];
        }
      }
      bar();
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
      for (int j = 0; j < NX; j += 4) {
        s0 += S.U[j];
        s1 += S.U[j + 1];
        s2 += S.U[j + 2];
        s3 += S.U[j + 3];
      }
      double c_temp = ((s0 + s1) + (s2 + s3)) / (double)NX;
      c_temp = dmax(c_temp, 1.0);  // ||q||_inf = 0 -> limit_scaling -> 1
      c_temp = limit_scaling(c_temp);
      cscale *= 1.0 / c_temp;
    }
  }
  bar();  // every thread has summed the cost terms before U is reused as factorisation workspace
  const double cinv = 1.0 / cscale;
  if (dbg) {
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      if (isvar) dbg[gvar(k, a)] = R.d[s];
      dbg[NX + grow(k, r1)] = R.e1[s];
      if (isrow2) dbg[NX + grow(k, 12 + l)] = S.E[44 * k + 12 + l];
    }
    if (tid == 0) dbg[NX + NR] = cscale;
  }
#pragma unroll
  for (int s = 0; s < SPW; ++s) R.b1[s] = R.e1[s] * R.b1[s];

  // ---- KKT factorisation
  RhoSet Rh(rho_now);
  int n_factor = 1;
  factor(C, S, R, Rh, cscale, tid);

  // ---- ADMM (osqp_solve)
  int status = ST_UNSOLVED;
  int iter = 0;
  for (iter = 1; iter <= C.max_iter; ++iter) {
    double tv[SPW], wv[SPW];
    // (A) u = E (rho z - y) per row; rhs = sigma x + D A0' u per variable
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      if (isrow2) {
        S.U[44 * k + l] = R.e1[s] * (Rh.rho_eq * R.z1[s] - R.y1[s]);
        S.U[44 * k + 12 + l] = S.E[44 * k + 12 + l] * (Rh.rho_eq * 0.0 - R.y2[s]);
      } else {
        const bool eq = (cc == 4) && (25.0 * R.e1[s] < 1e-4);
        S.U[44 * k + r1] = R.e1[s] * ((eq ? Rh.rho_eq : Rh.rho) * R.z1[s] - R.y1[s]);
      }
    }
    bar();
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      const double* u = S.U + 44 * k;
      double v = 0.0;
      if (isF) {
        v = dynT_force(C, S.Bang + 36 * k, u, a, ac);
        if ((sb[k] >> af) & 1) v += u[12 + a];
        const double* uc = u + 24 + 5 * af;
        if (ac == 0)
          v += uc[0] - uc[1];
        else if (ac == 1)
          v += uc[2] - uc[3];
        else
          v += -C.mu * (uc[0] + uc[1] + uc[2] + uc[3]) - uc[4];
      } else if (isvar) {
        v = -u[xi];
        if (k < NS - 1) {
          v += u[44 + xi];
          if (xi >= 6) v += C.dt * u[44 + xi - 6];
        }
      }
      tv[s] = C.sigma * R.x[s] + R.d[s] * v;
      if (isvar) S.TV[24 * k + a] = tv[s];
    }
    wsync();
    // (B) p^_{k+1} = Jx_{k+1} [S_k t_k]_X
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      double dv = 0.0;
      if (l >= 12 && l < 24) {
        const double* t = S.TV + 24 * k;
        const double* sf = S.SF + 288 * k + a;
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
        for (int c = 0; c < 12; c += 4) {
          a0 += sf[24 * c] * t[c];
          a1 += sf[24 * (c + 1)] * t[c + 1];
          a2 += sf[24 * (c + 2)] * t[c + 2];
          a3 += sf[24 * (c + 3)] * t[c + 3];
        }
#pragma unroll
        for (int c = 0; c < 12; c += 4) {
          a0 += R.sx[s][c] * t[12 + c];
          a1 += R.sx[s][c + 1] * t[13 + c];
          a2 += R.sx[s][c + 2] * t[14 + c];
          a3 += R.sx[s][c + 3] * t[15 + c];
        }
        dv = R.d[s] * ((a0 + a1) + (a2 + a3));
      }
      if (isvar) S.DX[24 * k + a] = dv;
    }
    wsync();
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      if (isrow2 && k < NS - 1) {
        const double* dx = S.DX + 24 * k + 12;
        double jw = dx[l];
        if (l < 6) jw += C.dt * dx[l + 6];
        S.PH[12 * (k + 1) + l] = S.E[44 * (k + 1) + l] * jw;
      }
    }
    bar();
    // (C) forward recurrence p_{k+1} = p^_{k+1} - T_k' p_k   (one warp)
    if (w == 0) {
      double p = (l < 12) ? S.PH[12 + l] : 0.0;
      for (int k = 1; k <= NS - 2; ++k) {
        const double* T = S.TT + 156 * (k - 1) + (l < 12 ? l : 0);
        double pj[12];
        gather12(p, pj);
        double a0 = (l < 12) ? S.PH[12 * (k + 1) + l] : 0.0, a1 = 0.0, a2 = 0.0;
#pragma unroll
        for (int j = 0; j < 12; j += 3) {
          a0 -= T[13 * j] * pj[j];
          a1 -= T[13 * (j + 1)] * pj[j + 1];
          a2 -= T[13 * (j + 2)] * pj[j + 2];
        }
        p = (a0 + a1) + a2;
        if (l < 12) S.PH[12 * (k + 1) + l] = p;
      }
    }
    bar();
    // (D) t' = t - rho G_k' p_k ; w = S_k t' ; g^_k = G_k w
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      if (isrow2) S.U[44 * k + l] = R.e1[s] * S.PH[12 * k + l];
    }
    wsync();
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      const double* pe = S.U + 44 * k;
      double v = 0.0;
      if (isF)
        v = dynT_force(C, S.Bang + 36 * k, pe, a, ac);
      else if (isvar)
        v = -pe[xi];
      if (isvar) S.TV[24 * k + a] = tv[s] - Rh.rho_eq * (R.d[s] * v);
    }
    wsync();
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      double acc = 0.0;
      if (isvar) {
        const double* t = S.TV + 24 * k;
        const double* sf = S.SF + 288 * k + a;
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
        for (int c = 0; c < 12; c += 4) {
          a0 += sf[24 * c] * t[c];
          a1 += sf[24 * (c + 1)] * t[c + 1];
          a2 += sf[24 * (c + 2)] * t[c + 2];
          a3 += sf[24 * (c + 3)] * t[c + 3];
        }
#pragma unroll
        for (int c = 0; c < 12; c += 4) {
          a0 += R.sx[s][c] * t[12 + c];
          a1 += R.sx[s][c + 1] * t[13 + c];
          a2 += R.sx[s][c + 2] * t[14 + c];
          a3 += R.sx[s][c + 3] * t[15 + c];
        }
        acc = (a0 + a1) + (a2 + a3);
        S.DX[24 * k + a] = R.d[s] * acc;
      }
      wv[s] = acc;
    }
    wsync();
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      if (isrow2 && k >= 1) S.GH[12 * k + l] = R.e1[s] * dyn_row_own(C, S.Bang + 36 * k, S.DX + 24 * k, l);
    }
    bar();
    // (E) backward recurrence g_k = g^_k - T_k g_{k+1}   (one warp)
    if (w == 1) {
      double g = (l < 12) ? S.GH[12 * (NS - 1) + l] : 0.0;
      for (int k = NS - 2; k >= 1; --k) {
        const double* T = S.TT + 156 * (k - 1) + 13 * (l < 12 ? l : 0);
        double gj[12];
        gather12(g, gj);
        double a0 = (l < 12) ? S.GH[12 * k + l] : 0.0, a1 = 0.0, a2 = 0.0;
#pragma unroll
        for (int j = 0; j < 12; j += 3) {
          a0 -= T[j] * gj[j];
          a1 -= T[j + 1] * gj[j + 1];
          a2 -= T[j + 2] * gj[j + 2];
        }
        g = (a0 + a1) + a2;
        if (l < 12) S.GH[12 * k + l] = g;
      }
    }
    bar();
    // (F) x~_k = w_k - S_k(:,X) u_k ,  u_k = rho D_X Jx_{k+1}' g_{k+1}
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      if (l >= 12 && l < 24) {
        double uj = 0.0;
        if (k < NS - 1) {
          const double* En = S.E + 44 * (k + 1);
          const double* g = S.GH + 12 * (k + 1);
          double acc = En[xi] * g[xi];
          if (xi >= 6) acc += C.dt * (En[xi - 6] * g[xi - 6]);
          uj = Rh.rho_eq * (R.d[s] * acc);
        }
        S.UX[12 * k + xi] = uj;
      }
    }
    wsync();
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      if (isvar) {
        const double* u = S.UX + 12 * k;
        double a0 = 0.0, a1 = 0.0, a2 = 0.0;
#pragma unroll
        for (int j = 0; j < 12; j += 3) {
          a0 += R.sx[s][j] * u[j];
          a1 += R.sx[s][j + 1] * u[j + 1];
          a2 += R.sx[s][j + 2] * u[j + 2];
        }
        wv[s] = wv[s] - ((a0 + a1) + a2);  // x~
        S.DX[24 * k + a] = R.d[s] * wv[s];
      }
    }
    bar();
#ifdef QRW_DEBUG_SOLVE
    if (dbg && iter == QRW_DEBUG_SOLVE) {
#pragma unroll
      for (int s = 0; s < SPW; ++s) {
        const int k = SPW * w + s;
        if (isvar) {
          dbg[gvar(k, a)] = tv[s];
          dbg[NX + gvar(k, a)] = wv[s];
        }
      }
    }
#endif
    // (G) z~ = A x~, relaxation, projection, dual update, x update
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      const int k = SPW * w + s;
      const double* dx = S.DX + 24 * k;
      if (isrow2) {
        const int i = l;
        double sm = -dx[12 + i];
        if (k > 0) {
          sm += dx[-24 + 12 + i];
          if (i < 6) sm += C.dt * dx[-24 + 18 + i];
        }
        if (i >= 6) {
          if (i < 9) {
            const int c = i - 6;
            sm += C.dtm * (dx[c] + dx[3 + c] + dx[6 + c] + dx[9 + c]);
          } else {
            const double* Br = S.Bang + 36 * k + 12 * (i - 9);
            double t = 0.0;
#pragma unroll
            for (int b = 0; b < 12; ++b) t += Br[b] * dx[b];
            sm += t;
          }
        }
        const double zt = R.e1[s] * sm;
        const double zr = C.alpha * zt + (1.0 - C.alpha) * R.z1[s];
        const double zn = R.b1[s];
        R.y1[s] += Rh.rho_eq * (zr - zn);
        R.z1[s] = zn;
        // force-enable row (k, a): l = u = 0
        const double zt2 = ((sb[k] >> af) & 1) ? S.E[44 * k + 12 + l] * dx[a] : 0.0;
        const double zr2 = C.alpha * zt2 + (1.0 - C.alpha) * 0.0;
        R.y2[s] += Rh.rho_eq * (zr2 - 0.0);
      } else {
        const bool eq = (cc == 4) && (25.0 * R.e1[s] < 1e-4);
        const double rr = eq ? Rh.rho_eq : Rh.rho, ri = eq ? Rh.rinv_eq : Rh.rinv;
        const double zt = R.e1[s] * cone_row(C, dx, cf, cc);
        const double zr = C.alpha * zt + (1.0 - C.alpha) * R.z1[s];
        double zn = zr + ri * R.y1[s];
        if (cc == 4) zn = dmax(zn, -25.0 * R.e1[s]);
        zn = dmin(zn, 0.0);
        R.y1[s] += rr * (zr - zn);
        R.z1[s] = zn;
      }
      if (isvar) R.x[s] = C.alpha * wv[s] + (1.0 - C.alpha) * R.x[s];
    }

    if (iter % C.check_every == 0) {
      // residuals (auxil.c compute_pri_res / compute_dua_res / compute_*_tol)
      bar();  // everyone is done reading DX / U of this iteration
#pragma unroll
      for (int s = 0; s < SPW; ++s) {
        const int k = SPW * w + s;
        if (isvar) S.DX[24 * k + a] = R.d[s] * R.x[s];
        S.U[44 * k + r1] = R.e1[s] * R.y1[s];
        if (isrow2) S.U[44 * k + 12 + l] = S.E[44 * k + 12 + l] * R.y2[s];
      }
      bar();
      double m[12];
#pragma unroll
      for (int q = 0; q < 12; ++q) m[q] = 0.0;
#pragma unroll
      for (int s = 0; s < SPW; ++s) {
        const int k = SPW * w + s;
        const double* dx = S.DX + 24 * k;
        auto row_acc = [&](double ax, double zi, double e) {
          const double einv = 1.0 / e, rs = ax - zi;
          m[0] = dmax(m[0], fabs(rs));
          m[1] = dmax(m[1], fabs(einv * rs));
          m[2] = dmax(m[2], fabs(zi));
          m[3] = dmax(m[3], fabs(einv * zi));
          m[4] = dmax(m[4], fabs(ax));
          m[5] = dmax(m[5], fabs(einv * ax));
        };
        if (isrow2) {
          const int i = l;
          double sm = -dx[12 + i];
          if (k > 0) {
            sm += dx[-24 + 12 + i];
            if (i < 6) sm += C.dt * dx[-24 + 18 + i];
          }
          if (i >= 6) {
            if (i < 9) {
              const int c = i - 6;
              sm += C.dtm * (dx[c] + dx[3 + c] + dx[6 + c] + dx[9 + c]);
            } else {
              const double* Br = S.Bang + 36 * k + 12 * (i - 9);
              double t = 0.0;
#pragma unroll
              for (int b = 0; b < 12; ++b) t += Br[b] * dx[b];
              sm += t;
            }
          }
          row_acc(R.e1[s] * sm, R.z1[s], R.e1[s]);
          const double e2 = S.E[44 * k + 12 + l];
          row_acc(((sb[k] >> af) & 1) ? e2 * dx[a] : 0.0, 0.0, e2);
        } else {
          row_acc(R.e1[s] * cone_row(C, dx, cf, cc), R.z1[s], R.e1[s]);
        }
        if (isvar) {
          const double* u = S.U + 44 * k;
          double v, p0;
          if (isF) {
            v = dynT_force(C, S.Bang + 36 * k, u, a, ac);
            if ((sb[k] >> af) & 1) v += u[12 + a];
            const double* uc = u + 24 + 5 * af;
            if (ac == 0)
              v += uc[0] - uc[1];
            else if (ac == 1)
              v += uc[2] - uc[3];
            else
              v += -C.mu * (uc[0] + uc[1] + uc[2] + uc[3]) - uc[4];
            p0 = C.wf;
          } else {
            v = -u[xi];
            if (k < NS - 1) {
              v += u[44 + xi];
              if (xi >= 6) v += C.dt * u[44 + xi - 6];
            }
            p0 = C.w[xi];
          }
          const double aty = R.d[s] * v;
          const double px = (cscale * p0 * R.d[s] * R.d[s]) * R.x[s];
          const double ds = px + aty, dinv = 1.0 / R.d[s];
          m[6] = dmax(m[6], fabs(ds));
          m[7] = dmax(m[7], fabs(dinv * ds));
          m[8] = dmax(m[8], fabs(aty));
          m[9] = dmax(m[9], fabs(dinv * aty));
          m[10] = dmax(m[10], fabs(px));
          m[11] = dmax(m[11], fabs(dinv * px));
        }
      }
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        m[q] = warp_max(m[q]);
        if (l == 0) S.red[12 * w + q] = m[q];
      }
      bar();
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        double v = S.red[q];
        for (int ww = 1; ww < NT / 32; ++ww) v = dmax(v, S.red[12 * ww + q]);
        m[q] = v;
      }
      const double pri_s = m[0], pri_u = m[1], nz_s = m[2], nz_u = m[3], nax_s = m[4], nax_u = m[5], dua_s = m[6],
                   dua_u = cinv * m[7], naty_s = m[8], naty_u = m[9], npx_s = m[10], npx_u = m[11];
      if (pri_u > 1e30 || dua_u > 1e30) {
        status = ST_NON_CVX;
        break;
      }
      const double eps_pri = C.eps_abs + C.eps_rel * dmax(nz_u, nax_u);
      const double eps_dua = C.eps_abs + C.eps_rel * (cinv * dmax(naty_u, npx_u));
      if (pri_u < eps_pri && dua_u < eps_dua) {
        status = ST_SOLVED;
        break;
      }
      if (iter == C.max_iter) {
        const double e1 = 10.0 * C.eps_abs + 10.0 * C.eps_rel * dmax(nz_u, nax_u);
        const double e2 = 10.0 * C.eps_abs + 10.0 * C.eps_rel * (cinv * dmax(naty_u, npx_u));
        status = (pri_u < e1 && dua_u < e2) ? ST_SOLVED_INACCURATE : ST_MAX_ITER;
      }
      if (iter % C.adapt_every == 0) {
        const double pr = pri_s / (dmax(nz_s, nax_s) + 1e-10);
        const double dr = dua_s / (dmax(naty_s, npx_s) + 1e-10);
        double est = Rh.rho * sqrt(pr / (dr + 1e-10));
        est = dmin(dmax(est, 1e-6), 1e6);
        if (est > Rh.rho * C.adapt_tol || est < Rh.rho / C.adapt_tol) {
          Rh = RhoSet(est);
          bar();
          factor(C, S, R, Rh, cscale, tid);
          ++n_factor;
        }
      }
    }
    bar();
  }
  if (iter > C.max_iter) iter = C.max_iter;
  if (status == ST_UNSOLVED) status = ST_MAX_ITER;

  // ---- retrieve_result (MPC.cpp:569-599) + state carried to the next call
  bar();
  const bool bad = (status == ST_NON_CVX);
  const double nanv = nan("");
#pragma unroll
  for (int s = 0; s < SPW; ++s) {
    const int k = SPW * w + s;
    if (isvar) {
      const double val = R.d[s] * R.x[s];
      if (isF)
        out[12 + a + 24 * k] = bad ? nanv : val;
      else
        out[xi + 24 * k] = bad ? nanv : (val + xref[xi + 12 * (1 + k)]);
      st_d[ST_X + gvar(k, a)] = bad ? 0.0 : R.x[s];
    }
    const int g1 = grow(k, r1);
    st_d[ST_Z + g1] = bad ? 0.0 : R.z1[s];
    st_d[ST_Y + g1] = bad ? 0.0 : R.y1[s];
    if (isrow2) {
      st_d[ST_Z + grow(k, 12 + l)] = 0.0;
      st_d[ST_Y + grow(k, 12 + l)] = bad ? 0.0 : R.y2[s];
    }
  }
  for (int i = tid; i < 576; i += NT) st_d[ST_BANG + i] = S.Bang[i];
  if (tid < 32) st_i[tid] = gait[tid];
  if (tid < NS) st_i[32 + tid] = sb[tid];
  if (tid == 0) {
    st_d[ST_RHO] = Rh.rho;
    info[0] = iter;
    info[1] = status;
    info[2] = n_factor;
    info[3] = overrun;
    if (rho_out) *rho_out = Rh.rho;
  }
  bar();
  (void)ictl;
}

}  // namespace cta
}  // namespace qrw
