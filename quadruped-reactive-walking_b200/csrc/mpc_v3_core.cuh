// mpc_v3_core.cuh -- one Solo12 centroidal-MPC solve by ONE 256-thread CTA: "force-first" KKT elimination with a
// NESTED-DISSECTION solve of the state system (round 2; replaces the sequential state chain of mpc_v2_core.cuh).
//
// Same contract, same OSQP iterate sequence (up to round-off) and same persistent state as the earlier kernels; it
// replaces the same reference code:
//   MPC::run  /root/reference/src/MPC.cpp:626-649  (construct_gait :686, construct_S :665, update_ML :418,
//   update_NK :469, call_solver :501 -> OSQP, retrieve_result :569).
//
// Why.  ncu on the v2.11 kernel (profiles/r02b_*): 52 % of an ADMM iteration was the block-tridiagonal state chain --
// 17 dependent 12x12 products walked by three warps while the other five sat at the CTA barrier (barrier stall 40 % of
// all samples, issue slots 36 % busy).  Latency work on that chain was exhausted (v2.0 .. v2.11); the dependency
// depth itself had to go.  The state system  S X = s  (S = Schur complement of the force blocks, block tridiagonal,
// 16 stages x 12) is now solved by nested dissection with DENSE precomputed operators:
//     separators  Sigma = stages {3, 7, 11};  interiors  P0 = {0,1,2}, P1 = {4,5,6}, P2 = {8,9,10}, P3 = {12..15}
//     L1   y      = A_p^-1 s_p                       (four dense inverses 36x36 / 48x48, one row per thread)
//     L1b  X_Sig  = R^-1 (s_Sig - E y)               (R = S_SigSig - E A^-1 E'; one operator [R^-1 | -R^-1 E] 36 x 108,
//                                                     a row split over four lanes)
//     L2   X_p    = y - Z_p X_Sig                    (Z_p = A_p^-1 E_p', 24 entries per row)
// Three dependent steps of 28..48 FMAs per thread with ALL eight warps busy instead of 17 steps on one warp: 13.8 k
// FMAs per iteration instead of 7 k, which the FP64 pipe (17 % busy before) has room for.  It is still an exact solve
// of the same KKT system; cond(S) after Ruiz equilibration is ~1e4 (measured on the workloads), the partitioned
// solve is as accurate as an LU solve (rel. error 2e-14).
//   * the dense inverses A_p^-1 stay in shared memory where the in-place Gauss-Jordan left them (rows padded to
//     38 / 50 doubles: conflict-free LDS.128); Z rows, the rows of [R^-1 | -R^-1 E] and the rows of F_k^-1 live in the
//     threads' TENSOR-MEMORY lanes (tcgen05.ld/st as a per-lane scratchpad, 840 B/cycle/SM, no interference with
//     LDS -- tools/tmem_bench.cu): 24 + 48 + 56 = 128 columns per thread, two warps per lane quadrant, 256 columns
//     per CTA, two CTAs per SM.  No tcgen05.mma anywhere: the tensor cores have no f64 kind, their memory is used.
//   * the rest is the v2 design: 16 lanes per horizon stage, lane j < 12 of stage k owns force f_k[j], state
//     X_{k+1}[j], dynamics row (k,j), force-enable row (k,j) and up to two friction-cone rows of its foot; iterates in
//     registers of their owner; everything else a stage owns in one shared-memory record; F_k (12x12) eliminated first,
//     GF = E B0 D_f F^-1 and M = GF D_f B0' E formed at factorisation time.
#pragma once
#include "mpc_common.cuh"

namespace qrw {
namespace nd {

constexpr int NT = 256;  // threads per CTA
constexpr int LD = 14;   // padded row length of the 6x12 GF block in the record (conflict-free LDS.128)

// ---- nested dissection of the 16 stages: stage k = 4 p + a; separator iff a == 3 and p < 3
constexpr int AP_LD0 = 38, AP_LD3 = 50;  // row strides of A_p^-1: 4 * odd words modulo 32 -> 8 consecutive rows tile the 32 banks
QRW_HD bool is_sep(int k) { return (k & 3) == 3 && k < 12; }
QRW_HD int ap_ld(int p) { return p < 3 ? AP_LD0 : AP_LD3; }
QRW_HD int ap_n(int p) { return p < 3 ? 36 : 48; }
QRW_HD int ap_off(int p) { return p * 36 * AP_LD0; }

// shared slab layout (doubles).  Everything a stage owns sits in one RECORD of SR doubles, so that a thread
// addresses all of it (and the records of the neighbouring stages) from one base register.
constexpr int V_AP = 0;                                 // A_p (during factorisation) / A_p^-1: 3 x [36][38] + [48][50]
constexpr int V_E = V_AP + 3 * 36 * AP_LD0 + 48 * AP_LD3;  // 6 coupling blocks [12][12]: 2 i = G_{s-1,s} ([m][c]), 2 i + 1 = G_{s,s+1} ([c][m])
constexpr int V_R = V_E + 6 * 144;                      // [36][36] R, then R^-1; dead after the factorisation: staging vectors
constexpr int V_SV = V_R;                               // [192] s_k of the interior stages (index 12 k + j)
constexpr int V_VEC = V_R + 192;                        // [112] s of the separators (36) | y of stages 2,4,6,8,10,12 (72) | 0 (4)
constexpr int V_XSP = V_VEC + 112;                      // [60]  0 (12) | X_3, X_7, X_11 (36) | 0 (12)
constexpr int V_REC = V_R + 36 * 36;                    // [16][SR] stage records
constexpr int R_BA = 0, R_BB = 16, R_BC = 32, R_BD = 48;  // 4 x 16 stage-internal exchange buffers
constexpr int R_XT = 88;   // [12] x~_X of the stage
constexpr int R_D = 112;   // [24] Ruiz column scaling: 0..11 force, 12..23 state
constexpr int R_E = 136;   // [44] Ruiz row scaling: 12 dyn | 12 enable | 20 cone
constexpr int R_BANG = 180;  // [3][12] angular block of B_k (unscaled)
constexpr int R_B1 = 216;  // [12] scaled right-hand side of the dynamics rows
constexpr int R_GF = 228;  // [6][14] rows 6..11 of GF = E B0 D_f F^-1 (what the dynamics rows see of a force right-hand side)
constexpr int R_MR = 312;  // [6][6]  M = GF D_f B0' E (rows 6..11 x rows 6..11)
constexpr int SR = 348;
// residual checks park two more vectors in the record (iteration-only slots, free at that point)
constexpr int R_A = 64;    // [12] residual check: E y of the dynamics rows
// tensor-memory columns of a thread (32-bit columns; 12 doubles = 24 columns): warps 0..3 use [0, 128), warps 4..7 [128, 256)
constexpr int TM_FI = 0;    // [24] row j of F_k^-1 of the thread
constexpr int TM_Z = 24;    // [48] row of Z_p of the thread: 12 entries for the separator below its partition, 12 for the one above
constexpr int TM_RE = 72;   // [56] 28 entries of a row of [R^-1 | -R^-1 E] (threads 0..143: row t / 4, entries 28 (t % 4) ..)
static_assert(TM_RE + 56 <= cta::TMEM_COLS / 2, "tensor-memory columns of a thread");
constexpr int V_RED = V_REC + 16 * SR;  // 128   reductions
constexpr int V_INT = V_RED + 128;      // 128 ints
constexpr int V3_SLAB_DOUBLES = V_INT + 64;
constexpr int V3_SLAB_BYTES = V3_SLAB_DOUBLES * 8;
// factorisation scratch inside a record, aliased on the iteration-only vectors [0, 112)
constexpr int R_Y = 0;   // [12][6] F_k^-1 (D_f B0')(:, rows 6..11); dead after step (c)
constexpr int R_N = 72;  // [36]    N_k (6x6); dead after step (d)
constexpr int R_PROW = 0, R_PCOL = 48;  // records 0..3 (A_p) and 4 (R): pivot row / column snapshot of the in-place Gauss-Jordan
static_assert(R_N + 36 <= R_D, "factorisation scratch overruns the record");
static_assert((SR * 8) % 16 == 0 && (V_REC * 8) % 16 == 0 && (V_VEC * 8) % 16 == 0 && (V_XSP * 8) % 16 == 0, "16-byte alignment");
static_assert(V_XSP + 60 <= V_REC, "staging vectors overrun the R slot");
static_assert(V3_SLAB_BYTES <= 233472 / 2 - 1024, "two CTAs per SM: 228 KB of shared memory, 1 KB reserved per CTA");
constexpr int V3_TMEM_SLOT = 120;  // index in the int buffer where tcgen05.alloc parks the CTA's tensor-memory base address

constexpr int V_PARK = V_E;  // [4][192] cone-row iterates (z, y of slot A and B) of thread 12 k + j between phases (1) and (7);
                             // the E slot is only live inside factor(), which the caller brackets with a reload / store
static_assert(4 * 192 <= 6 * 144, "cone-row park overruns the E slot");
struct Sm {
  double *AP, *E, *R, *SV, *Vec, *XSP, *rec, *red;
  int* ib;
  QRW_HD explicit Sm(double* b)
      : AP(b + V_AP), E(b + V_E), R(b + V_R), SV(b + V_SV), Vec(b + V_VEC), XSP(b + V_XSP), rec(b + V_REC), red(b + V_RED),
        ib(reinterpret_cast<int*>(b + V_INT)) {}
};

QRW_HD int grow(int k, int r) { return r < 12 ? 12 * k + r : (r < 24 ? 192 + 12 * k + (r - 12) : 384 + 20 * k + (r - 24)); }

// per-thread registers of variable lane j = 3 foot + axis of stage k (lanes 12..15 of a stage idle):
//   row j of F_k^-1 lives in tensor memory (TM_FI), row j of GF and column j of B_k in the record (R_GF, R_BANG)
//   xf, xX: scaled iterates of f_k[j], X_{k+1}[j]; z1, y1, e1: dynamics row (k,j); y2: force-enable row (k,j)
//   friction-cone rows of the foot, two slots: axis x -> rows 0,1 (+-fx - mu fz), axis y -> rows 2,3, axis z -> row 4 (-fz)
//   in slot A and nothing in slot B
struct Regs {
  double xf, xX, dF, dX, z1, y1, e1, y2;
  double zA, yA, eA, zB, yB, eB;
};

// unscaled dynamics row 6+r (r = 0..5) of a stage applied to a force-space vector v[0..11] held in shared memory
QRW_HD double b0_row(const MpcConst& C, const double* Bk, const double* v, int r) {
  if (r < 3) return C.dtm * (v[r] + v[3 + r] + v[6 + r] + v[9 + r]);
  const double* Br = Bk + 12 * (r - 3);
  double a0 = 0.0, a1 = 0.0, a2 = 0.0;
#pragma unroll
  for (int a = 0; a < 12; a += 3) {
    a0 += Br[a] * v[a];
    a1 += Br[a + 1] * v[a + 1];
    a2 += Br[a + 2] * v[a + 2];
  }
  return (a0 + a1) + a2;
}

QRW_HD double dot12(const double* m, const double* v) {
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
  for (int jj = 8; jj >= 0; jj -= 4) {
    a0 += m[jj] * v[jj];
    a1 += m[jj + 1] * v[jj + 1];
    a2 += m[jj + 2] * v[jj + 2];
    a3 += m[jj + 3] * v[jj + 3];
  }
  return (a0 + a1) + (a2 + a3);
}
// same product with the four partial sums accumulated in ascending order (the order phase (3) has always used)
QRW_HD double dot12a(const double* m, const double* v) {
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
  for (int b = 0; b < 12; b += 4) {
    a0 += m[b] * v[b];
    a1 += m[b + 1] * v[b + 1];
    a2 += m[b + 2] * v[b + 2];
    a3 += m[b + 3] * v[b + 3];
  }
  return (a0 + a1) + (a2 + a3);
}

// ------------------------------------------------------------------------------------------
// Factorisation of K, forces first.  All threads of the CTA; ends with a CTA barrier.
//   F_k   = K(f_k, f_k)                                   -> inverse: tensor memory (TM_FI); GF and M: record (R_GF, R_MR)
//   N_k   = e^2 (B0 D_f F_k^-1 D_f B0') e^2  (6 x 6)      -> record scratch
//   G_kk, G_{k,k+1}: Schur complement on the states       -> Sig / T slots, then the twisted block LDL'
//   (pivot inverses Sig_k, coupling matrices T_k = Sig_k G_{k,k+1} for k <= 7, U_k = Sig_k G_{k-1,k}' for k >= 9)
// ------------------------------------------------------------------------------------------
// A real call (noinline): inlined into the ADMM loop its register needs (a 28-entry operator segment, 12 + 12 accumulators)
// made ptxas spill the solver state in the hot phases of every iteration; as a call the spill code stays at the call site.
// (Everything by value: a reference to the kernel's parameter block would force a copy of it into local memory.)
struct FacIn {
  double dF, dX, e1;          // Ruiz scaling of the thread's force column, state column and dynamics row
  double wj;                  // state weight w[j] of the thread
  double dtm, mu, wf, sigma, dt;
};
QRW_HDN void factor(double* slab, const FacIn R, const RhoSet Rh, double cscale, int tid, unsigned tm) {
  const Sm S(slab);
  const FacIn& C = R;
  const int k = tid >> 4, j = tid & 15;
  const bool isv = j < 12;
  const int c3 = j % 3, ft = isv ? j / 3 : 0;
  const int* sb = S.ib + 32;
  double* rec = S.rec + SR * k;
  const double* Dk = rec + R_D;
  const double* Ek = rec + R_E;
  const double* Bk = rec + R_BANG;
  double* Y = rec + R_Y;
  double* N = rec + R_N;
  double F[12];
  // B_k(9..11, j) from the stage record (kept in registers across the ADMM loop they cost six registers for nothing)
  const double b0 = isv ? Bk[j] : 0.0, b1 = isv ? Bk[12 + j] : 0.0, b2 = isv ? Bk[24 + j] : 0.0;
  // ---- (a) row j of F_k
  if (isv) {
    const double e6 = Ek[6 + c3], e9 = Ek[9], e10 = Ek[10], e11 = Ek[11];
    const double* Ec = Ek + 24 + 5 * ft;
    const double c0 = Ec[0], c1 = Ec[1], c2 = Ec[2], c3e = Ec[3], c4 = Ec[4];
    const double r4 = (25.0 * c4 < 1e-4) ? Rh.rho_eq : Rh.rho;
    const double e2 = Ek[12 + j];
#pragma unroll
    for (int b = 0; b < 12; ++b) {
      double acc = (b % 3 == c3) ? (C.dtm * e6) * (C.dtm * e6) : 0.0;
      acc += (b0 * e9) * (Bk[b] * e9);
      acc += (b1 * e10) * (Bk[12 + b] * e10);
      acc += (b2 * e11) * (Bk[24 + b] * e11);
      const double dd = R.dF * Dk[b];
      double val = Rh.rho_eq * (acc * dd);
      if (b / 3 == ft) {
        const int cb = b % 3;
        const int lo = c3 < cb ? c3 : cb, hi = c3 < cb ? cb : c3;
        double cg = 0.0;
        if (lo == 0 && hi == 0)
          cg = Rh.rho * (c0 * c0 + c1 * c1);
        else if (lo == 1 && hi == 1)
          cg = Rh.rho * (c2 * c2 + c3e * c3e);
        else if (lo == 2)
          cg = Rh.rho * C.mu * C.mu * (c0 * c0 + c1 * c1 + c2 * c2 + c3e * c3e) + r4 * c4 * c4;
        else if (lo == 0 && hi == 2)
          cg = Rh.rho * C.mu * (c1 * c1 - c0 * c0);
        else if (lo == 1 && hi == 2)
          cg = Rh.rho * C.mu * (c3e * c3e - c2 * c2);
        val += cg * dd;
      }
      if (b == j) {
        val += cscale * C.wf * R.dF * R.dF + C.sigma;
        if ((sb[k] >> ft) & 1) {
          const double sd = e2 * R.dF;
          val += Rh.rho_eq * sd * sd;
        }
      }
      F[b] = val;
    }
  } else {
#pragma unroll
    for (int b = 0; b < 12; ++b) F[b] = (b == j) ? 1.0 : 0.0;
  }
  // ---- (b) Gauss-Jordan inverse inside the 16-lane group, row j in registers, pivot row by shuffle
#pragma unroll
  for (int p = 0; p < 12; ++p) {
    double pr[12];
#pragma unroll
    for (int c = 0; c < 12; ++c) pr[c] = cta::shfl16(F[c], p);
    const double d = 1.0 / pr[p];
    if (j == p) {
#pragma unroll
      for (int c = 0; c < 12; ++c) F[c] = (c == p) ? d : pr[c] * d;
    } else {
      const double f = F[p];
#pragma unroll
      for (int c = 0; c < 12; ++c) F[c] = (c == p) ? -f * d : F[c] - f * (pr[c] * d);
    }
  }
  // row j of F_k^-1 goes to the thread's own tensor-memory lane (warp-uniform: lanes 12..15 park their unit row)
  cta::tmem_st12(tm + 128 * (tid >> 7) + TM_FI, F);
  // ---- (c) Y(j, r) = sum_b Finv(j,b) D_f(b) B0(6+r, b), r = 0..5
  if (isv) {
    double y[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int b = 0; b < 12; ++b) {
      const double fd = F[b] * Dk[b];
      y[b % 3] += C.dtm * fd;
      y[3] += Bk[b] * fd;
      y[4] += Bk[12 + b] * fd;
      y[5] += Bk[24 + b] * fd;
    }
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      Y[6 * j + r] = y[r];
      const double g = Ek[6 + r] * y[r];  // GF(6+r, j) = E (B0 D_f F^-1)(6+r, j), F^-1 symmetric
      rec[R_GF + LD * r + j] = g;
    }
  }
  cta::wsync();
  // N_k(r, r') = e_r^2 e_r'^2 sum_a B0(6+r, a) D_f(a) Y(a, r'); three entries per variable lane
  double nv[3] = {0.0, 0.0, 0.0};
  if (isv) {
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      const int e = 3 * j + t, r = e / 6, rp = e - 6 * r;
      double acc = 0.0;
      if (r < 3) {
        for (int f = 0; f < 4; ++f) acc += C.dtm * (Dk[3 * f + r] * Y[6 * (3 * f + r) + rp]);
      } else {
        const double* Br = Bk + 12 * (r - 3);
        for (int a = 0; a < 12; ++a) acc += Br[a] * (Dk[a] * Y[6 * a + rp]);
      }
      const double er = Ek[6 + r], erp = Ek[6 + rp];
      nv[t] = (er * er) * acc * (erp * erp);
      rec[R_MR + e] = er * acc * erp;
    }
  }
  cta::wsync();
  if (isv) {
#pragma unroll
    for (int t = 0; t < 3; ++t) N[3 * j + t] = nv[t];
  }
  cta::bar();
  // ---- (d) rows of G_kk and G_{k,k+1}, written where the nested dissection wants them: interior blocks into the dense
  // A_p (both triangles), separator diagonal blocks into R, the six blocks next to a separator into E
  const int pk = k >> 2, ak = k & 3;
  const bool sepk = is_sep(k);
  if (isv) {
    const int i = j;
    const double* En = Ek + SR;  // dynamics-row scaling of stage k+1 (valid for k < 15)
    const double* Dn = Dk + SR;  // column scaling of stage k+1
    const double* Nn = N + SR;
    const double r2 = Rh.rho_eq * Rh.rho_eq;
    const int ld = ap_ld(pk);
    double* Ap = S.AP + ap_off(pk);
    const bool nsep = is_sep(k + 1);
#pragma unroll
    for (int ip = 0; ip < 12; ++ip) {
      double v = 0.0, o = 0.0;
      if (ip == i) {
        v = cscale * R.wj * R.dX * R.dX + C.sigma;
        const double g = R.e1 * R.dX;
        v += Rh.rho_eq * g * g;
        if (k < NS - 1) {
          const double a1 = En[i] * R.dX;
          v += Rh.rho_eq * a1 * a1;
          if (i >= 6) {
            const double a2 = C.dt * En[i - 6] * R.dX;
            v += Rh.rho_eq * a2 * a2;
          }
          o = -Rh.rho_eq * (En[i] * En[i]) * (R.dX * Dn[12 + i]);
        }
      }
      if (k < NS - 1) {
        const int lo = i < ip ? i : ip, hi = i < ip ? ip : i;
        if (hi - lo == 6) v += Rh.rho_eq * C.dt * En[lo] * En[lo] * Dk[12 + lo] * Dk[12 + hi];
        if (ip == i - 6) o = -Rh.rho_eq * C.dt * (En[ip] * En[ip]) * (R.dX * Dn[12 + ip]);
      }
      if (i >= 6 && ip >= 6) {
        const int e = 6 * (i - 6) + (ip - 6);
        double nn = N[e];
        if (k < NS - 1) {
          const double nx = Nn[e];
          nn += nx;
          o += r2 * (R.dX * Dn[12 + ip]) * nx;
        }
        v -= r2 * (R.dX * Dk[12 + ip]) * nn;
      }
      if (sepk) {
        S.R[36 * (12 * pk + i) + 12 * pk + ip] = v;
        S.E[144 * (2 * pk + 1) + 12 * i + ip] = o;  // G_{s,s+1}: [c = i][m = ip]
      } else {
        Ap[ld * (12 * ak + i) + 12 * ak + ip] = v;
        if (nsep) {
          S.E[144 * (2 * pk) + 12 * i + ip] = o;  // G_{s-1,s}: [m = i][c = ip]
        } else if (k < NS - 1) {
          Ap[ld * (12 * ak + i) + 12 * (ak + 1) + ip] = o;
          Ap[ld * (12 * (ak + 1) + ip) + 12 * ak + i] = o;
        }
      }
    }
    // blocks of A_p two stages apart are zero
    if (!sepk) {
      const int n = ap_n(pk) / 12;
      for (int a2 = 0; a2 < n; ++a2)
        if (a2 > ak + 1 || a2 < ak - 1)
#pragma unroll
          for (int ip = 0; ip < 12; ++ip) Ap[ld * (12 * ak + i) + 12 * a2 + ip] = 0.0;
    }
  }
  cta::bar();
  // ---- (e1) the four A_p -> A_p^-1 by Gauss-Jordan IN PLACE (no pivoting: SPD), all four at once.  Per pivot: the
  // scaled pivot row and the pivot column are snapshotted into the scratch of record p, then every entry is updated
  // from the snapshot.  Entries are dealt to the threads round robin (5..6 of each 36x36, 9 of the 48x48).
  for (int piv = 0; piv < 48; ++piv) {
    if (tid < 192) {
      const int p = tid / 48, c = tid - 48 * p;
      const int n = ap_n(p), ld = ap_ld(p);
      if (piv < n && c < n) {
        const double* Ap = S.AP + ap_off(p);
        double* sc = S.rec + SR * p;
        const double d = 1.0 / Ap[ld * piv + piv];
        sc[R_PROW + c] = (c == piv) ? d : Ap[ld * piv + c] * d;
        sc[R_PCOL + c] = Ap[ld * c + piv];
      }
    }
    cta::bar();
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const int n = ap_n(p), ld = ap_ld(p);
      if (piv >= n) continue;
      double* Ap = S.AP + ap_off(p);
      const double* prow = S.rec + SR * p + R_PROW;
      const double* pcol = S.rec + SR * p + R_PCOL;
      for (int e = tid; e < n * n; e += NT) {
        const int ei = e / n, ej = e - n * ei;
        double v;
        if (ei == piv)
          v = prow[ej];
        else if (ej == piv)
          v = -pcol[ei] * prow[piv];
        else
          v = Ap[ld * ei + ej] - pcol[ei] * prow[ej];
        Ap[ld * ei + ej] = v;
      }
    }
    cta::bar();
  }
  const unsigned tmc = tm + 128 * (tid >> 7);  // this thread's tensor-memory columns
  // ---- (e2) Z rows: Z(row, .) = A_p^-1(row, first stage) G_{s,s+1}' (separator below the partition) |
  //                               A_p^-1(row, last stage) G_{s-1,s} (separator above); warp-uniform tensor-memory stores
  {
    double zl[12], zr[12];
#pragma unroll
    for (int c = 0; c < 12; ++c) zl[c] = zr[c] = 0.0;
    if (isv && !sepk) {
      const double* Ar = S.AP + ap_off(pk) + ap_ld(pk) * (12 * ak + j);
      if (pk > 0) {
        double ar[12];
        cta::lds12(Ar, ar);
        const double* Eb = S.E + 144 * (2 * (pk - 1) + 1);  // [c][m]
#pragma unroll
        for (int c = 0; c < 12; ++c) zl[c] = dot12a(ar, Eb + 12 * c);
      }
      if (pk < 3) {
        double ar[12];
        cta::lds12(Ar + 24, ar);
        const double* Eb = S.E + 144 * (2 * pk);  // [m][c]
#pragma unroll
        for (int c = 0; c < 12; ++c) {
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int m = 0; m < 12; m += 2) {
            a0 += ar[m] * Eb[12 * m + c];
            a1 += ar[m + 1] * Eb[12 * m + 12 + c];
          }
          zr[c] = a0 + a1;
        }
      }
    }
    cta::tmem_st12(tmc + TM_Z, zl);
    cta::tmem_st12(tmc + TM_Z + 24, zr);
  }
  // ---- (e3) R = S_SigSig - E A^-1 E': thread t < 108 owns the 12 entries R(row t / 3, block t % 3)
  if (tid < 108) {
    const int r = tid / 3, sbk = tid - 3 * r, sa = r / 12, c = r - 12 * sa;
    double acc[12];
    double* Rr = S.R + 36 * r + 12 * sbk;
    if (sbk == sa) {
#pragma unroll
      for (int q = 0; q < 12; ++q) acc[q] = Rr[q];
    } else {
#pragma unroll
      for (int q = 0; q < 12; ++q) acc[q] = 0.0;
    }
    // one term: t(m') = sum_m g(m) A^-1(ra + m, ca + m'); acc(c') -= sum_m' t(m') h(m', c')
    auto term = [&](const double* g, int gs, const double* Ainv, int ld, const double* h, int hs_m, int hs_c) {
      double t[12];
#pragma unroll
      for (int mp = 0; mp < 12; ++mp) t[mp] = 0.0;
      for (int m = 0; m < 12; ++m) {
        const double gm = g[gs * m];
#pragma unroll
        for (int mp = 0; mp < 12; ++mp) t[mp] += gm * Ainv[ld * m + mp];
      }
#pragma unroll
      for (int cp = 0; cp < 12; ++cp) {
        double a = 0.0;
#pragma unroll
        for (int mp = 0; mp < 12; ++mp) a += t[mp] * h[hs_m * mp + hs_c * cp];
        acc[cp] -= a;
      }
    };
    const double* Elo = S.E + 144 * (2 * sa);      // G_{s-1,s}: [m][c]   (stage below the separator: last of partition sa)
    const double* Ehi = S.E + 144 * (2 * sa + 1);  // G_{s,s+1}: [c][m]   (stage above: first of partition sa + 1)
    const double* Alo = S.AP + ap_off(sa);
    const double* Ahi = S.AP + ap_off(sa + 1);
    const int ldlo = ap_ld(sa), ldhi = ap_ld(sa + 1);
    if (sbk == sa) {
      term(Elo + c, 12, Alo + ldlo * 24 + 24, ldlo, Elo, 12, 1);
      term(Ehi + 12 * c, 1, Ahi, ldhi, Ehi, 1, 12);
    } else if (sbk == sa + 1) {
      // through partition sa + 1: rows of its first stage, columns of its last stage; h = G_{s'-1,s'} of separator sbk: [m][c]
      term(Ehi + 12 * c, 1, Ahi + 24, ldhi, S.E + 144 * (2 * sbk), 12, 1);
    } else if (sbk == sa - 1) {
      // through partition sa: rows of its last stage, columns of its first stage; h = G_{s',s'+1} of separator sbk: [c][m]
      term(Elo + c, 12, Alo + ldlo * 24, ldlo, S.E + 144 * (2 * sbk + 1), 1, 12);
    }
#pragma unroll
    for (int q = 0; q < 12; ++q) Rr[q] = acc[q];  // nobody else reads or writes these twelve entries in this step
  }
  cta::bar();
  // ---- (e4) R -> R^-1 in place (36 pivots; snapshot in the scratch of record 4)
  for (int piv = 0; piv < 36; ++piv) {
    double* sc = S.rec + SR * 4;
    if (tid < 36) {
      const double d = 1.0 / S.R[36 * piv + piv];
      sc[R_PROW + tid] = (tid == piv) ? d : S.R[36 * piv + tid] * d;
      sc[R_PCOL + tid] = S.R[36 * tid + piv];
    }
    cta::bar();
    for (int e = tid; e < 1296; e += NT) {
      const int ei = e / 36, ej = e - 36 * ei;
      double v;
      if (ei == piv)
        v = sc[R_PROW + ej];
      else if (ej == piv)
        v = -sc[R_PCOL + ei] * sc[R_PROW + piv];
      else
        v = S.R[e] - sc[R_PCOL + ei] * sc[R_PROW + ej];
      S.R[e] = v;
    }
    cta::bar();
  }
  // ---- (e5) rows of [R^-1 | -R^-1 E] (36 x 108, padded to 112): thread t < 144 keeps entries 28 q .. 28 q + 27 of row
  // t / 4 in its tensor-memory columns.  Column 36 + 12 a6 + m belongs to y of stage 2, 4, 6, 8, 10, 12 (a6 = 0..5).
  {
    double seg[28];
#pragma unroll
    for (int i = 0; i < 28; ++i) seg[i] = 0.0;
    if (tid < 144) {
      const int r = tid >> 2, q = tid & 3;
      const double* Rr = S.R + 36 * r;
#pragma unroll
      for (int i = 0; i < 28; ++i) {
        const int v = 28 * q + i;
        if (v < 36) {
          seg[i] = Rr[v];
        } else if (v < 108) {
          const int a6 = (v - 36) / 12, m = (v - 36) - 12 * a6, si = a6 >> 1;
          const double* Eb = S.E + 144 * a6;  // a6 even: G_{s-1,s} [m][c]; odd: G_{s,s+1} [c][m]
          const int sm = (a6 & 1) ? 1 : 12, sc = (a6 & 1) ? 12 : 1;
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int c = 0; c < 12; c += 2) {
            a0 += Rr[12 * si + c] * Eb[sm * m + sc * c];
            a1 += Rr[12 * si + c + 1] * Eb[sm * m + sc * (c + 1)];
          }
          seg[i] = -(a0 + a1);
        }
      }
    }
    cta::tmem_st12(tmc + TM_RE, seg);
    cta::tmem_st12(tmc + TM_RE + 24, seg + 12);
    cta::tmem_st4(tmc + TM_RE + 48, seg + 24);
  }
  cta::tmem_wait_st();
  cta::bar();  // R is dead: its slot holds the staging vectors of the iterations from here on; their padding is zero
  if (tid < 12) S.XSP[tid] = 0.0, S.XSP[48 + tid] = 0.0;
  if (tid < 4) S.Vec[108 + tid] = 0.0;
  cta::bar();
}

// ------------------------------------------------------------------------------------------
// One full MPC::run for one instance; every thread of the CTA calls it with the same arguments.
// Same I/O contract as mpc_solve_instance (mpc_core.cuh).  tmem = base address of the CTA's tensor-memory allocation
// (cta::TMEM_COLS columns).
// ------------------------------------------------------------------------------------------
// The eight instance pointers are derived from the batch base pointers (kernel parameters: constant bank) and the instance
// index where they are used -- at the head and, from an opaque copy of the index, again at the tail -- so that they do not
// occupy sixteen registers across the ADMM loop.
struct MpcIo {
  const double *xref, *fs;
  double* st_d;
  int* st_i;
  double* out;
  int* info;
  double *rho, *dbg;
};
QRW_HD void mpc_v3_solve(const MpcConst& C, double* slab, const MpcIo& io, int b, int first, unsigned tmem) {
  const double* xref = io.xref + (size_t)b * 12 * (NS + 1);
  const double* fs = io.fs + (size_t)b * 12 * C.N_gait;
  double* st_d = io.st_d + (size_t)b * MPC_STATE_DOUBLES;
  int* st_i = io.st_i + (size_t)b * MPC_STATE_INTS;
  double* dbg = io.dbg ? io.dbg + (size_t)b * 1281 : nullptr;
  const Sm S(slab);
  const unsigned tm = cta::tmem_addr(tmem, 0);  // this thread's lane quadrant
  const int tid = cta::tid();
  const int k = tid >> 4, j = tid & 15, w = tid >> 5, l = tid & 31;
  const unsigned tmc = tm + 128 * (w >> 2);     // this thread's 128 tensor-memory columns
  const int pk = k >> 2, ak = k & 3;            // partition and position of the stage in the nested dissection
  const bool sepk = is_sep(k);
  const int NG = C.N_gait;
  int* gait = S.ib;       // [32]
  int* sb = S.ib + 32;    // [16] S_gait bits per step (bit f set = foot f swings, force pinned to 0)
  int* itmp = S.ib + 48;  // [64]
  Regs R;
  const bool isv = j < 12;
  const int c3 = j % 3, ft = isv ? j / 3 : 0;
  // friction-cone rows owned by this lane (stage-local row index; slot B only for the x and y axes)
  const int rowA = 24 + 5 * ft + (c3 < 2 ? 2 * c3 : 4);
  const bool hasB = isv && c3 < 2;
  int rec_off = SR * k;
#if QRW_ON_GPU
  // opaque to the compiler: otherwise ptxas re-derives the record address from threadIdx (S2R + 4 integer ops, ~45 cycles
  // of latency) at the head of every phase of the ADMM loop instead of keeping one register live (SASS of r01b)
  asm volatile("" : "+r"(rec_off));
#endif
  double* rec = S.rec + rec_off;
  double* BAk = rec + R_BA;
  double* BBk = rec + R_BB;
  double* BCk = rec + R_BC;
  double* BDk = rec + R_BD;
  double* Bk = rec + R_BANG;
  double* Dk = rec + R_D;
  double* Ek = rec + R_E;

  // ---- load the state carried between calls
  for (int i = tid; i < 576; i += NT) S.rec[SR * (i / 36) + R_BANG + (i % 36)] = first ? 0.0 : st_d[ST_BANG + i];
  if (tid < 32) gait[tid] = first ? 0 : st_i[tid];
  if (tid < NS) sb[tid] = first ? 0 : st_i[32 + tid];
  R.xf = R.xX = R.z1 = R.y1 = R.y2 = R.zA = R.yA = R.zB = R.yB = 0.0;
  if (!first && isv) {
    R.xf = st_d[ST_X + gvar(k, j)];
    R.xX = st_d[ST_X + gvar(k, 12 + j)];
    R.z1 = st_d[ST_Z + grow(k, j)];
    R.y1 = st_d[ST_Y + grow(k, j)];
    R.y2 = st_d[ST_Y + grow(k, 12 + j)];
    R.zA = st_d[ST_Z + grow(k, rowA)];
    R.yA = st_d[ST_Y + grow(k, rowA)];
    if (hasB) {
      R.zB = st_d[ST_Z + grow(k, rowA + 1)];
      R.yB = st_d[ST_Y + grow(k, rowA + 1)];
    }
  }
  double rho_now = first ? C.rho0 : st_d[ST_RHO];

  // ---- construct_gait (MPC.cpp:686-701), construct_S (:665-681): warp 0
  if (w == 0) {
    int zero = 1, bits = 0;
    if (l < NG) {
      for (int c = 0; c < 12; ++c) {
        const double v = fs[l + NG * c];
        if (!(fabs(v) <= 1e-12)) zero = 0;
      }
      for (int f = 0; f < 4; ++f)
        if (fs[l + NG * (3 * f)] != 0.0) bits |= (1 << f);
    }
    itmp[l] = zero;
    itmp[32 + l] = bits;
  }
  cta::bar();
  int nact = 0;
  while (nact < NG && !itmp[nact]) ++nact;
  const int overrun = (nact >= NG);
  cta::bar();
  if (w == 0) {
    if (l < nact)
      gait[l] = itmp[32 + l];
    else if (l == nact && l < NG)
      gait[l] = 0;
  }
  cta::bar();
  int nS = 0;
  while (nS < NG && gait[nS] != 0) ++nS;
  cta::bar();
  if (tid < nS && tid < NS) sb[tid] = (~gait[tid]) & 15;
  // ---- B_k angular blocks (create_ML :213-232 on the first call, update_ML :418-452 after)
  if (tid < NS && (first || tid < nS)) {
    const int kk = tid;
    const double yaw = xref[5 + 12 * kk];
    const double cy = cos(yaw), sy = sin(yaw);
    const double Rm[3][3] = {{cy, -sy, 0.0}, {sy, cy, 0.0}, {0.0, 0.0, 1.0}};
    double T[3][3], M[3][3], Ii[3][3];
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) {
        double acc = 0.0;
        for (int q = 0; q < 3; ++q) acc += Rm[q][a] * C.gI[3 * q + b];
        T[a][b] = acc;
      }
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) {
        double acc = 0.0;
        for (int q = 0; q < 3; ++q) acc += T[a][q] * Rm[q][b];
        M[a][b] = acc;
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
    for (int f = 0; f < 4; ++f) {
      double v[3];
      for (int r = 0; r < 3; ++r)
        v[r] = first ? (C.foot[4 * r + f] - xref[r + 12 * kk])
                     : (fs[kk + NG * (3 * f + r)] - (xref[r + 12 * kk] + C.off_com[r]));
      const double Sk[3][3] = {{0.0, -v[2], v[1]}, {v[2], 0.0, -v[0]}, {-v[1], v[0], 0.0}};
      for (int r = 0; r < 3; ++r)
        for (int c2 = 0; c2 < 3; ++c2) {
          double acc = 0.0;
          for (int q = 0; q < 3; ++q) acc += Ii[r][q] * Sk[q][c2];
          S.rec[SR * kk + R_BANG + 12 * r + 3 * f + c2] = C.dt * acc;
        }
    }
  }
  // ---- right-hand side of the dynamics rows (create_NK :261-289 / update_NK :469-486), row owners
  double bnd = 0.0;
  if (isv) {
    const int i = j;
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
  // Ruiz start values
  double e2 = 1.0;
  R.dF = R.dX = R.e1 = R.eA = R.eB = 1.0;
  if (isv) {
    Dk[j] = 1.0;
    Dk[12 + j] = 1.0;
    Ek[j] = 1.0;
    Ek[12 + j] = 1.0;
    Ek[rowA] = 1.0;
    if (hasB) Ek[rowA + 1] = 1.0;
  }
  cta::bar();
  double bc0 = 0.0, bc1 = 0.0, bc2 = 0.0;  // B_k(9..11, j), live during the Ruiz passes only
  if (isv) bc0 = Bk[j], bc1 = Bk[12 + j], bc2 = Bk[24 + j];
  const int bit = isv ? ((sb[k] >> ft) & 1) : 0;

  // ---- Ruiz equilibration, 10 passes (OSQP scale_data; rerun on every call, MPC.cpp:541,548)
  double cscale = 1.0;
  for (int pass = 0; pass < 10; ++pass) {
    double dDF = 1.0, dDX = 1.0, dE1 = 1.0, dE2 = 1.0, dEA = 1.0, dEB = 1.0;
    if (isv) {
      const int i = j;
      // column of f_k[j]
      const double d = R.dF;
      double nrm = cscale * C.wf * d * d;
      nrm = dmax(nrm, C.dtm * Ek[6 + c3] * d);
      nrm = dmax(nrm, fabs(bc0) * Ek[9] * d);
      nrm = dmax(nrm, fabs(bc1) * Ek[10] * d);
      nrm = dmax(nrm, fabs(bc2) * Ek[11] * d);
      if (bit) nrm = dmax(nrm, e2 * d);
      const double* Ec = Ek + 24 + 5 * ft;
      if (c3 < 2) {
        nrm = dmax(nrm, dmax(R.eA, R.eB) * d);
      } else {
        const double em = dmax(dmax(Ec[0], Ec[1]), dmax(Ec[2], Ec[3]));
        nrm = dmax(nrm, C.mu * em * d);
        nrm = dmax(nrm, R.eA * d);
      }
      dDF = 1.0 / sqrt(limit_scaling(nrm));
      // column of X_{k+1}[i]
      const double dx = R.dX;
      double nx = fabs(cscale * C.w[i] * dx * dx);
      nx = dmax(nx, R.e1 * dx);
      if (k < NS - 1) {
        nx = dmax(nx, Ek[SR + i] * dx);
        if (i >= 6) nx = dmax(nx, C.dt * Ek[SR + i - 6] * dx);
      }
      dDX = 1.0 / sqrt(limit_scaling(nx));
      // dynamics row (k, i)
      const double e = R.e1;
      double nr = e * dx;
      if (k > 0) {
        nr = dmax(nr, e * Dk[-SR + 12 + i]);
        if (i < 6) nr = dmax(nr, C.dt * e * Dk[-SR + 18 + i]);
      }
      if (i >= 6) {
        if (i < 9) {
          for (int f = 0; f < 4; ++f) nr = dmax(nr, C.dtm * e * Dk[3 * f + i - 6]);
        } else {
          const double* Br = Bk + 12 * (i - 9);
          for (int b = 0; b < 12; ++b) nr = dmax(nr, fabs(Br[b]) * e * Dk[b]);
        }
      }
      dE1 = 1.0 / sqrt(limit_scaling(nr));
      const double nf = bit ? e2 * d : 0.0;
      dE2 = 1.0 / sqrt(limit_scaling(nf));
      // friction-cone rows of this lane
      const double dz = Dk[3 * ft + 2];
      if (c3 < 2) {
        dEA = 1.0 / sqrt(limit_scaling(dmax(R.eA * d, C.mu * R.eA * dz)));
        dEB = 1.0 / sqrt(limit_scaling(dmax(R.eB * d, C.mu * R.eB * dz)));
      } else {
        dEA = 1.0 / sqrt(limit_scaling(R.eA * d));
      }
    }
    cta::bar();
    double psum = 0.0;
    if (isv) {
      R.dF *= dDF;
      R.dX *= dDX;
      R.e1 *= dE1;
      e2 *= dE2;
      R.eA *= dEA;
      R.eB *= dEB;
      Dk[j] = R.dF;
      Dk[12 + j] = R.dX;
      Ek[j] = R.e1;
      Ek[12 + j] = e2;
      Ek[rowA] = R.eA;
      if (hasB) Ek[rowA + 1] = R.eB;
      psum = fabs(cscale * C.wf * R.dF * R.dF) + fabs(cscale * C.w[j] * R.dX * R.dX);
    }
    // cost normalisation: mean of the column norms of the scaled P (diagonal)
#pragma unroll
    for (int m = 8; m > 0; m >>= 1) psum += cta::shfl_xor(psum, m);
    if (j == 0) S.red[k] = psum;
    cta::bar();
    double tot = 0.0;
#pragma unroll
    for (int kk = 0; kk < NS; ++kk) tot += S.red[kk];
    double c_temp = tot / (double)NX;
    c_temp = dmax(c_temp, 1.0);  // ||q||_inf = 0 -> limit_scaling -> 1
    c_temp = limit_scaling(c_temp);
    cscale *= 1.0 / c_temp;
  }
  if (tid == 0) S.red[120] = cscale;  // parked for factor() and the residual checks; published by the barrier below
  cta::bar();
  if (dbg && isv) {
    dbg[gvar(k, j)] = R.dF;
    dbg[gvar(k, 12 + j)] = R.dX;
    dbg[NX + grow(k, j)] = R.e1;
    dbg[NX + grow(k, 12 + j)] = e2;
    dbg[NX + grow(k, rowA)] = R.eA;
    if (hasB) dbg[NX + grow(k, rowA + 1)] = R.eB;
    if (tid == 0) dbg[NX + NR] = cscale;
  }
  if (isv) rec[R_B1 + j] = R.e1 * bnd;
  // the axis-z lane's row 4 is the only cone row with a finite lower bound (-25 E); OSQP makes it an equality
  // row (rho_eq) when the scaled width drops under RHO_TOL
  const bool eq4 = isv && c3 == 2 && (25.0 * R.eA < 1e-4);

  // ---- KKT factorisation (inside the loop: first iteration and after every rho update)
  RhoSet Rh(rho_now);
  // The cone-row iterates are touched in phases (1) and (7) only: between them they sit in shared memory (the E slot of the
  // factorisation), not in registers -- with them the ADMM loop needed ~12 registers more than it has and ptxas spilled to
  // local memory, which at 2 x 113 KB of shared memory per SM has almost no L1 behind it (long-scoreboard stalls, r02d).
  double* cpk = S.E + (12 * k + (isv ? j : 0));
  double czA = R.zA, cyA = R.yA, czB = R.zB, cyB = R.yB;
  if (tid == 0) S.red[121] = Rh.rinv, S.red[122] = Rh.rinv_eq;
  int n_factor = 0;
  bool refactor = true;

  // ---- ADMM (osqp_solve)
  int status = ST_UNSOLVED;
  int iter = 0;
  int until_check = C.check_every;  // countdown instead of iter % check_every (an integer division per iteration)
  for (iter = 1; iter <= C.max_iter; ++iter) {
    if (refactor) {
      if (iter > 1 && isv) czA = cpk[0], cyA = cpk[192], czB = cpk[384], cyB = cpk[576];
      factor(slab, FacIn{R.dF, R.dX, R.e1, C.w[j < 12 ? j : 0], C.dtm, C.mu, C.wf, C.sigma, C.dt}, Rh, S.red[120], tid, tm);
      if (isv) cpk[0] = czA, cpk[192] = cyA, cpk[384] = czB, cpk[576] = cyB;
      ++n_factor;
      refactor = false;
    }
    const double rhoA = eq4 ? Rh.rho_eq : Rh.rho;
    // (1) row values u = E (rho z - y); pair sums of the cone rows for the axis-z lane of the foot
    double u1 = 0.0, u2 = 0.0, uA = 0.0, uB = 0.0;
    if (isv) {
      u1 = R.e1 * (Rh.rho_eq * R.z1 - R.y1);
      if (bit) u2 = Ek[12 + j] * (Rh.rho_eq * 0.0 - R.y2);
      const double eA = Ek[rowA], eB = hasB ? Ek[rowA + 1] : 1.0;
      uA = eA * (rhoA * cpk[0] - cpk[192]);
      uB = eB * (Rh.rho * cpk[384] - cpk[576]);
      BAk[j] = u1;
      BDk[j] = uA + uB;
    }
    cta::wsync();
    // row j of F_k^-1 from the thread's tensor-memory lane: requested here, needed in phase (3)
    cta::TRow tfi;
    cta::tmem_ld12(tmc + TM_FI, tfi);
    // (2) t_f = sigma x_f + D_f (A' u)_f
    if (isv) {
      double v = C.dtm * BAk[6 + c3];
      v += Bk[j] * BAk[9];
      v += Bk[12 + j] * BAk[10];
      v += Bk[24 + j] * BAk[11];
      if (bit) v += u2;
      if (c3 < 2)
        v += uA - uB;
      else
        v += -C.mu * (BDk[3 * ft] + BDk[3 * ft + 1]) - uA;
      BBk[j] = C.sigma * R.xf + R.dF * v;
    }
    cta::wsync();
    // (3) u_f = F^-1 t_f and w = GF t_f (rows 6..11) from the same broadcast of t_f;
    //     m = u - rho E w: what the state system sees of the dynamics rows; A_k = sigma x_X - D_X m_k
    double uf = 0.0, wv = 0.0, mk = 0.0;
    double fv[12];
    cta::tmem_get12(tfi, fv);  // warp-uniform (tcgen05.wait::ld is warp-collective): outside the lane branch
    if (isv) {
      // row of GF as six LDS.128, issued back to back with the broadcast of t_f: left to itself ptxas loaded it one
      // LDS.64 at a time right in front of the FMA that uses it (a shared-memory latency per product)
      double tv[12];
      cta::lds12(BBk, tv);
      double gv[12];  // lanes j < 6 have no GF row: they read the (valid, aligned) words in front of it and drop the product
      cta::lds12(rec + R_GF + LD * (j - 6), gv);
      uf = dot12a(fv, tv);
      const double wq = dot12a(gv, tv);
      if (j >= 6) wv = wq;
      mk = u1 - Rh.rho_eq * (R.e1 * wv);
      BAk[j] = mk;
    }
    cta::bar();
    // (4) right-hand side of the state system, formed by the owner of the entry:
    //     s_k = (sigma x_X - D_X m_k) + D_X A' m_{k+1}; interior stages into SV, separators into the head of Vec
    if (isv) {
      double sv = C.sigma * R.xX - R.dX * mk;
      if (k < NS - 1) {
        double v = rec[SR + R_BA + j];
        if (j >= 6) v += C.dt * rec[SR + R_BA + j - 6];
        sv += R.dX * v;
      }
      if (sepk)
        S.Vec[12 * pk + j] = sv;
      else
        S.SV[12 * k + j] = sv;
    }
    // the two warps of a partition (stages 4 p .. 4 p + 3): s_p is complete.  Immediate barrier ids: with a run-time id
    // ptxas reserves all 16 hardware barriers for the CTA and only one CTA fits an SM
    if (pk == 0) cta::nb_sync(1, 64);
    else if (pk == 1) cta::nb_sync(2, 64);
    else if (pk == 2) cta::nb_sync(3, 64);
    else cta::nb_sync(4, 64);
    // (L1) y = A_p^-1 s_p: one row per thread, 36 (48) products; the rows come from shared memory (padded: conflict-free
    // LDS.128), the vector is a broadcast
    double yv = 0.0;
    if (isv && !sepk) {
      const double* Ar = S.AP + ap_off(pk) + ap_ld(pk) * (12 * ak + j);
      const double* sp = S.SV + 48 * pk;
      const int nch = pk == 3 ? 4 : 3;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 1
      for (int ch = 0; ch < nch; ++ch) {
        double m0[12], v0[12];
        cta::lds12(Ar + 12 * ch, m0);
        cta::lds12(sp + 12 * ch, v0);
#pragma unroll
        for (int b = 0; b < 12; b += 4) {
          a0 += m0[b] * v0[b];
          a1 += m0[b + 1] * v0[b + 1];
          a2 += m0[b + 2] * v0[b + 2];
          a3 += m0[b + 3] * v0[b + 3];
        }
      }
      yv = (a0 + a1) + (a2 + a3);
      // y of the stages next to a separator: stage 2, 4, 6, 8, 10, 12 -> slot 0..5
      if (ak == 2 && pk < 3) S.Vec[36 + 12 * (2 * pk) + j] = yv;
      if (ak == 0 && pk > 0) S.Vec[36 + 12 * (2 * pk - 1) + j] = yv;
    }
    // (L1b) X_Sig = [R^-1 | -R^-1 E] [s_Sig; y]: threads 0..143, row t / 4, 28 products per lane, summed over the lane quad.
    // The operator segment comes from tensor memory in three pieces, one in flight at a time (registers).
    cta::TRow ta;
    cta::TRow4 tc;
    cta::tmem_ld12(tmc + TM_RE, ta);  // every warp (a request under `if (w < 5)` makes the row look loop-carried to ptxas: 24 registers)
    cta::bar();
    if (w < 5) {
      // one operator piece in registers at a time: the next tensor-memory request goes out behind the product of the
      // previous piece (two pieces in flight cost 24 registers more than the loop has: ptxas spilled the solver state)
      double m[12], v[12];
      const double* vp = S.Vec + 28 * (l & 3);
      cta::lds12(vp, v);
      cta::tmem_get12(ta, m);
      double acc = dot12(m, v);
      cta::tmem_ld12(tmc + TM_RE + 24, ta);
      cta::lds12(vp + 12, v);
      cta::tmem_get12(ta, m);
      acc += dot12(m, v);
      cta::tmem_ld4(tmc + TM_RE + 48, tc);
      cta::lds4(vp + 24, v);
      cta::tmem_get4(tc, m);
      acc += (m[0] * v[0] + m[1] * v[1]) + (m[2] * v[2] + m[3] * v[3]);
      acc += cta::shfl_xor(acc, 1);
      acc += cta::shfl_xor(acc, 2);
      if ((l & 3) == 0 && tid < 144) {
        const int r = tid >> 2, si = r / 12, c = r - 12 * si;
        S.XSP[12 + r] = acc;
        S.rec[SR * (4 * si + 3) + R_XT + c] = acc;
      }
    }
    // (L2) X_p = y - Z_p X_Sig: 24 products per thread, the Z row from tensor memory (two pieces)
    cta::tmem_ld12(tmc + TM_Z, ta);
    cta::bar();
    {
      double z[12], x[12];
      cta::lds12(S.XSP + 12 * pk, x);
      cta::tmem_get12(ta, z);
      cta::tmem_ld12(tmc + TM_Z + 24, ta);
      const double d0 = dot12(z, x);
      cta::lds12(S.XSP + 12 * pk + 12, x);
      cta::tmem_get12(ta, z);
      const double d1 = dot12(z, x);
      if (isv && !sepk) rec[R_XT + j] = yv - (d0 + d1);
    }
    cta::bar();
    // (5) dynamics rows applied to the new states: r = E (-D_X x~_X + A D_X x~_X^{k-1})
    double xtX = 0.0, rX = 0.0;
    if (isv) {
      xtX = rec[R_XT + j];
      double smX = -(R.dX * xtX);
      if (k > 0) {
        smX += Dk[-SR + 12 + j] * rec[-SR + R_XT + j];
        if (j < 6) smX += C.dt * (Dk[-SR + 18 + j] * rec[-SR + R_XT + j + 6]);
      }
      rX = R.e1 * smX;
      BAk[j] = rX;
    }
    cta::wsync();
    // (6) x~_f = u_f - rho GF' r ;  z~ (dynamics rows) = r + w - rho M r
    double xtf = 0.0, zt = rX;
    if (isv) {
      // everything the phase reads is requested first (three LDS.128 for r, six LDS.64 for the GF column, three
      // LDS.128 for the M row); lanes j < 6 have no M row: they read the (valid, aligned) words in front of it and drop them
      double r[6], mr[6], gcv[6];
      cta::lds6(BAk + 6, r);
      const double* gc = rec + R_GF + j;  // column j of GF
#pragma unroll
      for (int q = 0; q < 6; ++q) gcv[q] = gc[q * LD];
      cta::lds6(rec + R_MR + 6 * (j - 6), mr);
      const double s0 = (gcv[0] * r[0] + gcv[1] * r[1]) + (gcv[2] * r[2] + gcv[3] * r[3]) + (gcv[4] * r[4] + gcv[5] * r[5]);
      xtf = uf - Rh.rho_eq * s0;
      BCk[j] = R.dF * xtf;
      const double s1 = (mr[0] * r[0] + mr[1] * r[1]) + (mr[2] * r[2] + mr[3] * r[3]) + (mr[4] * r[4] + mr[5] * r[5]);
      if (j >= 6) zt = rX + (wv - Rh.rho_eq * s1);
    }
    cta::wsync();
    // (7) relaxation, projection, dual update, x update
    if (isv) {
      const double zr = C.alpha * zt + (1.0 - C.alpha) * R.z1;
      const double zn = rec[R_B1 + j];
      R.y1 += Rh.rho_eq * (zr - zn);
      R.z1 = zn;
      // force-enable row (k, j): l = u = 0
      const double dxf = R.dF * xtf;
      const double zt2 = bit ? Ek[12 + j] * dxf : 0.0;
      const double zr2 = C.alpha * zt2 + (1.0 - C.alpha) * 0.0;
      R.y2 += Rh.rho_eq * (zr2 - 0.0);
      R.xf = C.alpha * xtf + (1.0 - C.alpha) * R.xf;
      R.xX = C.alpha * xtX + (1.0 - C.alpha) * R.xX;
      // friction-cone rows: slot A = +f_axis - mu fz (or -fz for the z lane), slot B = -f_axis - mu fz
      const double fz = BCk[3 * ft + 2];
      const double rowa = (c3 < 2) ? dxf - C.mu * fz : -fz;
      {
        const double eA = Ek[rowA], zA = cpk[0], yA = cpk[192], rinvA = S.red[eq4 ? 122 : 121];
        const double ztc = eA * rowa;
        const double zrc = C.alpha * ztc + (1.0 - C.alpha) * zA;
        double znc = zrc + rinvA * yA;
        if (c3 == 2) znc = dmax(znc, -25.0 * eA);
        znc = dmin(znc, 0.0);
        cpk[192] = yA + rhoA * (zrc - znc);
        cpk[0] = znc;
      }
      if (c3 < 2) {
        const double eB = Ek[rowA + 1], zB = cpk[384], yB = cpk[576];
        const double ztc = eB * (-dxf - C.mu * fz);
        const double zrc = C.alpha * ztc + (1.0 - C.alpha) * zB;
        const double znc = dmin(zrc + S.red[121] * yB, 0.0);
        cpk[576] = yB + Rh.rho * (zrc - znc);
        cpk[384] = znc;
      }
    }
    if (--until_check == 0) {
      until_check = C.check_every;
      // residuals (auxil.c compute_pri_res / compute_dua_res / compute_*_tol)
      cta::bar();  // everyone is done with the exchange buffers of this iteration
      double yA = 0.0, yB = 0.0;
      if (isv) {
        BCk[j] = R.dF * R.xf;
        rec[R_XT + j] = R.dX * R.xX;
        rec[R_A + j] = R.e1 * R.y1;
        BBk[j] = Ek[12 + j] * R.y2;
        yA = Ek[rowA] * cpk[192];
        yB = (hasB ? Ek[rowA + 1] : 1.0) * cpk[576];
        BDk[j] = yA + yB;
      }
      cta::bar();
      const double cscale = S.red[120], cinv = 1.0 / cscale;
      double m[12];
#pragma unroll
      for (int q = 0; q < 12; ++q) m[q] = 0.0;
      auto row_acc = [&](double ax, double zi, double e) {
        const double einv = 1.0 / e, rs = ax - zi;
        m[0] = dmax(m[0], fabs(rs));
        m[1] = dmax(m[1], fabs(einv * rs));
        m[2] = dmax(m[2], fabs(zi));
        m[3] = dmax(m[3], fabs(einv * zi));
        m[4] = dmax(m[4], fabs(ax));
        m[5] = dmax(m[5], fabs(einv * ax));
      };
      auto var_acc = [&](double aty, double px, double d) {
        const double ds = px + aty, dinv = 1.0 / d;
        m[6] = dmax(m[6], fabs(ds));
        m[7] = dmax(m[7], fabs(dinv * ds));
        m[8] = dmax(m[8], fabs(aty));
        m[9] = dmax(m[9], fabs(dinv * aty));
        m[10] = dmax(m[10], fabs(px));
        m[11] = dmax(m[11], fabs(dinv * px));
      };
      if (isv) {
        double sm = -rec[R_XT + j];
        if (k > 0) {
          sm += rec[-SR + R_XT + j];
          if (j < 6) sm += C.dt * rec[-SR + R_XT + j + 6];
        }
        if (j >= 6) sm += b0_row(C, Bk, BCk, j - 6);
        row_acc(R.e1 * sm, R.z1, R.e1);
        row_acc(bit ? Ek[12 + j] * BCk[j] : 0.0, 0.0, Ek[12 + j]);
        const double dxf = BCk[j], fz = BCk[3 * ft + 2];
        const double eA = Ek[rowA];
        row_acc(eA * ((c3 < 2) ? dxf - C.mu * fz : -fz), cpk[0], eA);
        if (c3 < 2) {
          const double eB = Ek[rowA + 1];
          row_acc(eB * (-dxf - C.mu * fz), cpk[384], eB);
        }
        // force variable
        const double* u = rec + R_A;
        double v = C.dtm * u[6 + c3];
        v += Bk[j] * u[9];
        v += Bk[12 + j] * u[10];
        v += Bk[24 + j] * u[11];
        if (bit) v += BBk[j];
        if (c3 < 2)
          v += yA - yB;
        else
          v += -C.mu * (BDk[3 * ft] + BDk[3 * ft + 1]) - yA;
        var_acc(R.dF * v, (cscale * C.wf * R.dF * R.dF) * R.xf, R.dF);
        // state variable
        double vx = -u[j];
        if (k < NS - 1) {
          vx += u[SR + j];
          if (j >= 6) vx += C.dt * u[SR + j - 6];
        }
        var_acc(R.dX * vx, (cscale * C.w[j] * R.dX * R.dX) * R.xX, R.dX);
      }
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        m[q] = cta::warp_max(m[q]);
        if (l == 0) S.red[16 + 12 * w + q] = m[q];
      }
      cta::bar();
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        double v = S.red[16 + q];
        for (int ww = 1; ww < NT / 32; ++ww) v = dmax(v, S.red[16 + 12 * ww + q]);
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
        const double t1 = 10.0 * C.eps_abs + 10.0 * C.eps_rel * dmax(nz_u, nax_u);
        const double t2 = 10.0 * C.eps_abs + 10.0 * C.eps_rel * (cinv * dmax(naty_u, npx_u));
        status = (pri_u < t1 && dua_u < t2) ? ST_SOLVED_INACCURATE : ST_MAX_ITER;
      }
      if (iter % C.adapt_every == 0) {
        const double pr = pri_s / (dmax(nz_s, nax_s) + 1e-10);
        const double dr = dua_s / (dmax(naty_s, npx_s) + 1e-10);
        double est = Rh.rho * sqrt(pr / (dr + 1e-10));
        est = dmin(dmax(est, 1e-6), 1e6);
        if (est > Rh.rho * C.adapt_tol || est < Rh.rho / C.adapt_tol) {
          Rh = RhoSet(est);
          if (tid == 0) S.red[121] = Rh.rinv, S.red[122] = Rh.rinv_eq;  // read in phase (7), behind the barriers of factor()
          refactor = true;
        }
      }
      cta::bar();
    }
  }
  if (iter > C.max_iter) iter = C.max_iter;
  if (status == ST_UNSOLVED) status = ST_MAX_ITER;

  // ---- retrieve_result (MPC.cpp:569-599) + state carried to the next call
  cta::bar();
  cta::opaque(b);
  xref = io.xref + (size_t)b * 12 * (NS + 1);
  st_d = io.st_d + (size_t)b * MPC_STATE_DOUBLES;
  st_i = io.st_i + (size_t)b * MPC_STATE_INTS;
  double* out = io.out + (size_t)b * 24 * NS;
  int* info = io.info + 4 * b;
  double* rho_out = io.rho ? io.rho + b : nullptr;
  const bool bad = (status == ST_NON_CVX);
  const double nanv = nan("");
  if (isv) {
    out[12 + j + 24 * k] = bad ? nanv : R.dF * R.xf;
    out[j + 24 * k] = bad ? nanv : (R.dX * R.xX + xref[j + 12 * (1 + k)]);
    st_d[ST_X + gvar(k, j)] = bad ? 0.0 : R.xf;
    st_d[ST_X + gvar(k, 12 + j)] = bad ? 0.0 : R.xX;
    st_d[ST_Z + grow(k, j)] = bad ? 0.0 : R.z1;
    st_d[ST_Y + grow(k, j)] = bad ? 0.0 : R.y1;
    st_d[ST_Z + grow(k, 12 + j)] = 0.0;
    st_d[ST_Y + grow(k, 12 + j)] = bad ? 0.0 : R.y2;
    st_d[ST_Z + grow(k, rowA)] = bad ? 0.0 : cpk[0];
    st_d[ST_Y + grow(k, rowA)] = bad ? 0.0 : cpk[192];
    if (hasB) {
      st_d[ST_Z + grow(k, rowA + 1)] = bad ? 0.0 : cpk[384];
      st_d[ST_Y + grow(k, rowA + 1)] = bad ? 0.0 : cpk[576];
    }
  }
  for (int i = tid; i < 576; i += NT) st_d[ST_BANG + i] = S.rec[SR * (i / 36) + R_BANG + (i % 36)];
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
  cta::bar();
}

}  // namespace nd
}  // namespace qrw
