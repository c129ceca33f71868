// mpc_v4_core.cuh -- one Solo12 centroidal-MPC solve by ONE 256-thread CTA: "force-first" KKT elimination (as
// mpc_v2_core.cuh) with the state chain folded onto SUPER-STAGES of two horizon stages (round 2).
//
// Same contract, same OSQP iterate sequence (up to round-off) and same persistent state as the earlier kernels; it
// replaces the same reference code:
//   MPC::run  /root/reference/src/MPC.cpp:626-649  (construct_gait :686, construct_S :665, update_ML :418,
//   update_NK :469, call_solver :501 -> OSQP, retrieve_result :569).
//
// Why.  In mpc_v2_core.cuh the state system (block tridiagonal, 16 stages x 12) is solved by a twisted chain of 17
// DEPENDENT 12x12 products per ADMM iteration; ncu (profiles/r02b_*) has that chain at 52 % of an iteration with five of
// the eight warps parked at the barrier, and v2.0 .. v2.11 had exhausted the per-step latency work.  A nested-dissection
// solve with dense operators (mpc_v3_core.cuh) removed the chain but paid for it with three barrier-separated levels that
// cost as much.  Here the chain stays inside ONE warp per direction (warp-level synchronisation only), but its depth halves:
//   super-stage m = stages (2 m, 2 m + 1), 24 states; the 8 x 8 block system is factored twisted (super-stages 0..3
//   downwards, 7..5 upwards, super-stage 4 last).  The coupling block between two super-stages has ONE non-zero 12 x 12
//   corner (stage 2 m + 1 <-> stage 2 m + 2), so T_m = Sigma_m C_m has 12 non-zero columns, and
//     elimination   s'_{m+1}[stage 2m+2] = s_{2m+2} - T_m[0:12,:]' s'_{2m} - T_m[12:24,:]' s_{2m+1}
//     back-subst.   X_{2m} = v_{2m} - T_m[0:12,:] X_{2m+2},   X_{2m+1} = v_{2m+1} - T_m[12:24,:] X_{2m+2}
//   of which only ONE 12 x 12 product per step is on the dependent path (s_{2m+1} is an input; X_{2m+1} is not needed by
//   the next step): 4 + 1 + 4 dependent 12x12 steps instead of 7 + 2 + 8.  The off-path products ride in the same warp in
//   the shadow of the dependent step's latencies; v_m = Sigma_m s'_m (24 x 24) is computed by two helper warps behind the
//   elimination.  13.8 k FMAs per iteration for the chain instead of 7 k -- the FP64 pipe was 17 % busy.
//   Four chain warps (elimination, two helpers, back-substitution) in the four lane quadrants of tensor memory, each with
//   96 doubles of matrix rows per lane (tcgen05.ld as a per-lane scratchpad, as in v2).
#pragma once
#include "mpc_common.cuh"

namespace qrw {
namespace v4 {

constexpr int NT = 256;  // threads per CTA
constexpr int LD = 14;   // padded row length of the 6x12 GF block in the record (conflict-free LDS.128)
constexpr int LS = 26;   // padded row length of a 24x24 super-stage block (52 words: 8 consecutive rows tile the 32 banks)
constexpr int SB = 24 * LS;

// shared slab layout (doubles).  Everything a stage owns sits in one RECORD of SR doubles, so that a thread
// addresses all of it (and the records of the neighbouring stages) from one base register.
constexpr int V_SD = 0;                // [8][24][LS] super-stage pivot blocks D_m, then their inverses Sigma_m
constexpr int V_GC = V_SD + 8 * SB;    // [7][12][12] coupling blocks G_{2m+1,2m+2} (row: stage 2m+1, column: stage 2m+2)
constexpr int V_TT = V_GC + 7 * 144;   // [7][24][12] T_m = Sigma_m C_m (m <= 3: slot m) | U_m = Sigma_m C_{m-1}' (m >= 5: slot m - 1)
constexpr int V_W = V_TT + 7 * 288;    // [24][24] Gauss-Jordan ping-pong
constexpr int V_REC = V_W + 576;       // [16][SR] stage records
constexpr int R_BA = 0, R_BB = 16, R_BC = 32, R_BD = 48;  // 4 x 16 stage-internal exchange buffers
constexpr int R_A = 64;    // [12] sigma x_X - D_X m_k            (written by the stage's own lanes); later v_k
constexpr int R_CC = 76;   // [12] D_X A' m_{k+1}                 (written by the lanes of stage k+1)
constexpr int R_XT = 88;   // [12] x~_X of the stage
constexpr int R_SP = 100;  // [12] forward-recurrence result s'_k
constexpr int R_D = 112;   // [24] Ruiz column scaling: 0..11 force, 12..23 state
constexpr int R_E = 136;   // [44] Ruiz row scaling: 12 dyn | 12 enable | 20 cone
constexpr int R_BANG = 180;  // [3][12] angular block of B_k (unscaled)
constexpr int R_B1 = 216;  // [12] scaled right-hand side of the dynamics rows
constexpr int R_GF = 228;  // [6][14] rows 6..11 of GF = E B0 D_f F^-1 (what the dynamics rows see of a force right-hand side)
constexpr int R_MR = 312;  // [6][6]  M = GF D_f B0' E (rows 6..11 x rows 6..11)
constexpr int SR = 348;
// tensor-memory columns of a lane quadrant (32-bit columns; a row of 12 doubles = 24 columns)
constexpr int TM_FI = 0;     // [2][24] row j of F_k^-1 of the thread: warps 0..3 at column 0, warps 4..7 at column 24
constexpr int TM_ROLE = 48;  // role rows of the quadrant: [0, 96) first role warp (E0 / B0), [96, 192) second (E1 / B1): 4 slots x 12
                             // doubles per lane; the helper warps H1 / H2 use all 192 columns (8 slots)
static_assert(TM_ROLE + 192 <= cta::TMEM_COLS, "tensor-memory columns of a CTA");
constexpr int V_RED = V_REC + 16 * SR;  // 128   reductions
constexpr int V_INT = V_RED + 128;      // 128 ints
constexpr int V4_SLAB_DOUBLES = V_INT + 64;
constexpr int V4_SLAB_BYTES = V4_SLAB_DOUBLES * 8;
// The four chain warps park their iterates around the chain call.  The park area aliases factorisation scratch that is dead
// between factorisations: the pivot blocks of super-stages 0..3 (super-stage 4 stays: the middle step reads Sigma_4 from it).
constexpr int PK = 18;                 // 16 doubles per lane, padded so that the lanes' 16-byte accesses fall on distinct banks
constexpr int V_PARK = V_SD;           // roles 0..3: [4][32][PK] on the blocks of super-stages 0..3; roles 4, 5: on those of 5, 6
QRW_HD int park_off(int role) { return role < 4 ? role * 32 * PK : 5 * SB + (role - 4) * 32 * PK; }
static_assert(4 * 32 * PK <= 4 * SB && 2 * 32 * PK <= 2 * SB, "the park area overruns the pivot blocks it aliases");
// factorisation scratch inside a record, aliased on the iteration-only vectors [0, 112)
constexpr int R_Y = 0;   // [12][6] F_k^-1 (D_f B0')(:, rows 6..11); dead after step (c)
constexpr int R_N = 72;  // [36]    N_k (6x6); dead after step (d)
static_assert(R_N + 36 <= R_D, "factorisation scratch overruns the record");
static_assert((SR * 8) % 16 == 0 && (V_REC * 8) % 16 == 0 && (SB * 8) % 16 == 0 && (LS * 8) % 16 == 0, "16-byte alignment");
static_assert(V4_SLAB_BYTES <= 233472 / 2 - 1024, "two CTAs per SM: 228 KB of shared memory, 1 KB reserved per CTA");
constexpr int V4_TMEM_SLOT = 120;  // index in the int buffer where tcgen05.alloc parks the CTA's tensor-memory base address

struct Sm {
  double *SD, *GC, *TT, *rec, *W, *red, *park;
  int* ib;
  QRW_HD explicit Sm(double* b)
      : SD(b + V_SD), GC(b + V_GC), TT(b + V_TT), rec(b + V_REC), W(b + V_W), red(b + V_RED), park(b + V_PARK),
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

// ------------------------------------------------------------------------------------------
// Factorisation of K, forces first.  All threads of the CTA; ends with a CTA barrier.
//   F_k   = K(f_k, f_k)                                   -> inverse: tensor memory (TM_FI); GF and M: record (R_GF, R_MR)
//   N_k   = e^2 (B0 D_f F_k^-1 D_f B0') e^2  (6 x 6)      -> record scratch
//   G_kk, G_{k,k+1}: Schur complement on the states       -> super-stage pivot blocks D_m and coupling blocks, then the
//   twisted block LDL' over the super-stages (Sigma_m, T_m / U_m)
// ------------------------------------------------------------------------------------------
QRW_HD void factor(const MpcConst& C, const Sm& S, Regs& R, const RhoSet Rh, double cscale, int tid, int recw, unsigned tm) {
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
  cta::tmem_st12(tm + TM_FI + 24 * (tid >> 7), F);
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
  // ---- (d) rows of G_kk and G_{k,k+1}, written into the super-stage layout: stage k = 2 m + h sits in rows / columns
  // 12 h .. 12 h + 11 of the pivot block D_m; the coupling of an even stage to the next one is the off-diagonal block of D_m
  // (both triangles), the coupling of an odd stage to the next super-stage goes to GC[m]
  const int sm_ = k >> 1, sh = k & 1;
  if (isv) {
    const int i = j;
    const double* En = Ek + SR;  // dynamics-row scaling of stage k+1 (valid for k < 15)
    const double* Dn = Dk + SR;  // column scaling of stage k+1
    const double* Nn = N + SR;
    const double r2 = Rh.rho_eq * Rh.rho_eq;
    double* Dm = S.SD + SB * sm_;
#pragma unroll
    for (int ip = 0; ip < 12; ++ip) {
      double v = 0.0, o = 0.0;
      if (ip == i) {
        v = cscale * C.w[i] * R.dX * R.dX + C.sigma;
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
      Dm[LS * (12 * sh + i) + 12 * sh + ip] = v;
      if (sh == 0) {
        Dm[LS * i + 12 + ip] = o;
        Dm[LS * (12 + ip) + i] = o;
      } else if (k < NS - 1) {
        S.GC[144 * sm_ + 12 * i + ip] = o;
      }
    }
  }
  cta::bar();
  // ---- (e) TWISTED block LDL' over the eight super-stages: 7, 6, 5 upwards, 0, 1, 2, 3 downwards, 4 last (with both Schur
  // updates).  Per super-stage: Gauss-Jordan inverse of the 24x24 pivot block (ping-pong D_m <-> W, 24 pivots: the result is
  // back in D_m), the 24x12 coupling product T_m = Sigma_m[:, 12:24] G (downwards) / U_m = Sigma_m[:, 0:12] G' (upwards), and
  // the Schur update of the 12x12 corner of the next pivot block.
  for (int step = 0; step < 8; ++step) {
    const int m = step < 3 ? 7 - step : (step < 7 ? step - 3 : 4);
    const bool upper = step < 3;
    double* G = S.SD + SB * m;
    for (int p = 0; p < 24; ++p) {
      const double* Wc = (p & 1) ? S.W : G;
      double* Wn = (p & 1) ? G : S.W;
      const int ldc = (p & 1) ? 24 : LS, ldn = (p & 1) ? LS : 24;
      const double d = 1.0 / Wc[ldc * p + p];
      for (int e = tid; e < 576; e += NT) {
        const int ei = e / 24, ej = e - 24 * ei;
        double v;
        if (ei == p)
          v = (ej == p) ? d : Wc[ldc * p + ej] * d;
        else if (ej == p)
          v = -Wc[ldc * ei + p] * d;
        else
          v = Wc[ldc * ei + ej] - Wc[ldc * ei + p] * (Wc[ldc * p + ej] * d);
        Wn[ldn * ei + ej] = v;
      }
      cta::bar();
    }
    if (m == 4) break;
    const int gslot = upper ? m - 1 : m;  // coupling block between this super-stage and the one that receives the update
    const double* Gc = S.GC + 144 * gslot;
    double* Tm = S.TT + 288 * gslot;
    for (int e = tid; e < 288; e += NT) {
      const int a = e / 12, c = e - 12 * a;
      double acc = 0.0;
#pragma unroll
      for (int r = 0; r < 12; ++r) acc += upper ? G[LS * a + r] * Gc[12 * c + r] : G[LS * a + 12 + r] * Gc[12 * r + c];
      Tm[e] = acc;
    }
    cta::bar();
    if (tid < 144) {
      const int c = tid / 12, cp = tid - 12 * c;
      double acc = 0.0;
      if (upper) {
#pragma unroll
        for (int r = 0; r < 12; ++r) acc += Gc[12 * c + r] * Tm[12 * r + cp];
        S.SD[SB * (m - 1) + LS * (12 + c) + 12 + cp] -= acc;
      } else {
#pragma unroll
        for (int r = 0; r < 12; ++r) acc += Gc[12 * r + c] * Tm[12 * (12 + r) + cp];
        S.SD[SB * (m + 1) + LS * c + cp] -= acc;
      }
    }
    cta::bar();
  }
  // ---- (f) the six chain warps copy what they walk through in every iteration into their tensor-memory lanes, 12 doubles
  // per slot.  Roles (warp recw + r): 0 = E0, 4 = E1 elimination of the lower / upper half, 1 = B0, 5 = B1 back-substitution,
  // 2 = H1, 3 = H2 helpers.  In E and B warps lanes 0..11 hold what is on the dependent path, lanes 16..27 the product next
  // to it (E: one step AHEAD, see state_elim).  Lanes 12..15 of a half-warp duplicate lane 11.
  const int w = tid >> 5, role = (w - recw) & 7;
  if (role < 6) {
    const int lane = tid & 31, sp = lane >> 4, i0 = lane & 15, i = i0 < 12 ? i0 : 11;
    const bool up = role >= 4;                       // E1 / B1: the upper half
    const unsigned tb = tm + TM_ROLE + (up ? 96 : 0);
    if (role == 0 || role == 4) {
      for (int t = 0; t < 4; ++t) {
        double c0[12];
        // dependent lanes: step t; the lanes next to them: step t + 1.  T_t sits in TT slot t, U_{7-t} in slot 6 - t.
        const int st = t + sp, nst = up ? 3 : 4;
        const bool live = st < nst;
        const double* Tm = S.TT + 288 * (live ? (up ? 6 - st : st) : 0) + i;
        const int r0 = (up ? !sp : sp) ? 12 : 0;  // lower: dependent rows 0..11, side rows 12..23; upper: the other way round
        for (int r = 0; r < 12; ++r) c0[r] = live ? Tm[12 * (r0 + r)] : 0.0;
        cta::tmem_st12(tb + 24 * t, c0);
      }
    } else if (role == 1 || role == 5) {
      for (int u = 0; u < 4; ++u) {
        double c0[12];
        const bool live = u < (up ? 3 : 4);
        // T_{3-u} (slot 3 - u): dependent row i, side row 12 + i  |  U_{5+u} (slot 4 + u): dependent row 12 + i, side row i
        const double* Tm = S.TT + 288 * (live ? (up ? 4 + u : 3 - u) : 0) + 12 * (((up ? !sp : sp) ? 12 : 0) + i);
        for (int c = 0; c < 12; ++c) c0[c] = live ? Tm[c] : 0.0;
        cta::tmem_st12(tb + 24 * u, c0);
      }
    } else {
      // helpers: row i of the first (role 2) / second (role 3) stage of Sigma_t (lanes 0..11) | Sigma_{7-t} (lanes 16..27)
      for (int t = 0; t < 4; ++t) {
        double c0[12], c1[12];
        const bool live = !(sp && t == 3);
        const double* Sr = S.SD + SB * (live ? (sp ? 7 - t : t) : 0) + LS * (12 * (role - 2) + i);
        cta::lds12(Sr, c0);
        cta::lds12(Sr + 12, c1);
        cta::tmem_st12(tm + TM_ROLE + 48 * t, c0);
        cta::tmem_st12(tm + TM_ROLE + 48 * t + 24, c1);
      }
    }
  }
  cta::tmem_wait_st();
  // the scratch overlapped CC of the last stage, which nobody writes and the recurrence reads as zero
  if (tid < 12) S.rec[SR * (NS - 1) + R_CC + tid] = 0.0;
  cta::bar();  // the park area aliases pivot blocks of super-stages other than 4: nobody reads them after this point
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
// c - m.v with c folded into the first partial sum: one addition less on the dependent path of a chain step
QRW_HD double sub_dot12(double c, const double* m, const double* v) {
  double a0 = -c, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
  for (int jj = 8; jj >= 0; jj -= 4) {
    a0 += m[jj] * v[jj];
    a1 += m[jj + 1] * v[jj + 1];
    a2 += m[jj + 2] * v[jj + 2];
    a3 += m[jj + 3] * v[jj + 3];
  }
  return -(a0 + a1) - (a2 + a3);
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
// The state system per ADMM iteration, SIX warps.  The lower half (super-stages 0..3, stages 0..7; k0 = 0, dk = +1) and the
// upper half (super-stages 7..5, stages 15..10; k0 = 15, dk = -1) each have an elimination warp and a back-substitution warp;
// super-stage 4 (stages 8, 9) is the middle.  Step t of an elimination reads s' of stage crit = k0 + 2 t dk (on the dependent
// path) and s of stage crit + dk (an input) and produces s' of stage out = crit + 2 dk:
//     s'_out = s_out - T[side rows]' s_side - T[dependent rows]' s'_crit
// Lanes 16..27 of the warp form acc = s_out - T[side rows]' s_side ONE STEP AHEAD and leave it in the R_SP slot of the stage;
// lanes 0..11 take it from there and subtract the dependent product: both half-warps run the same instruction stream (one
// 12 x 12 product per lane and step, as a step of mpc_v2_core.cuh), only the addresses differ.
//   E0 / E1 (warps recw, recw + 4)      elimination, then the middle: X_8 (E0) / X_9 (E1) = Sigma_4[row, :] [s''_8; s''_9]
//   H1 / H2 (warps recw + 2, recw + 3)  v of the first / second stage of each super-stage: v = Sigma_m[rows, :] s'_m
//   B0 / B1 (warps recw + 1, recw + 5)  back-substitution from X_8 / X_9 outwards; lanes 0..11 the stage the next step needs,
//                                       lanes 16..27 the stage next to it
// named barriers: 1, 2 (E0, E1 -> H1, H2: 128), 3 (H1, H2 -> B0, B1: 128), 4 (E0 -> B0: 64), 5 (E0 <-> E1: 64), 6 (E1 -> B1: 64).
// s' is passed through R_SP, v through R_A, X through R_XT of the stage records, as in mpc_v2_core.cuh.
// ------------------------------------------------------------------------------------------
template <int UP>
QRW_HDN void state_elim(const double* Sig4, const double* TTside, double* rec, int lane, unsigned tmr) {
  constexpr int K0 = (NS - 1) * UP, DK = 1 - 2 * UP, NST = 4 - UP;
  const int i0 = lane & 15;
  int sp = lane >> 4;  // 0: dependent path, 1: the product next to it
  cta::opaque(sp);     // keeps ptxas from cloning the chain per half-warp
  const int i = i0 < 12 ? i0 : 11;
  const bool wr = i0 < 12;
  // per-lane bases: everything below addresses a stage as base + (compile-time stage) * SR
  const double* bv = rec + R_SP + (sp ? 3 * DK * SR : 0);  // s'_crit of this step | s of the side stage of the NEXT step
  double* bs = rec + i + (sp ? 2 * DK * SR : 0);            // stage produced by this step | by the next one
  cta::TRow tq[2];
  cta::tmem_ld12(tmr, tq[0]);
  if (!UP) {
    // s of the odd stages 1, 3, 5, 7 (the even ones and stage 15 come ready from phases (3) / (4))
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      double* r = rec + SR * (1 + 4 * q) + (sp ? 2 * SR : 0);
      const double p = r[R_A + i] + r[R_CC + i];
      if (wr) r[R_SP + i] = p;
    }
    cta::wsync();
  }
  // acc of step 0 (lanes 16..27): the side rows of T_0 | U_7 from shared memory (column i)
  {
    double ts[12], pvs[12];
    double* ro = rec + SR * (K0 + 2 * DK);
    const double sn = UP ? ro[R_A + i] + ro[R_CC + i] : ro[R_SP + i];
    cta::lds12(rec + SR * (K0 + DK) + R_SP, pvs);
    cta::lds12c<12>(TTside + i, ts);
    const double a = sub_dot12(sn, ts, pvs);
    if (wr && sp) ro[R_SP + i] = a;
  }
#pragma unroll
  for (int t = 0; t < NST; ++t) {
    constexpr int dummy = 0;
    (void)dummy;
    const int kc = K0 + 2 * t * DK, ko = kc + 2 * DK;
    double pv[12], tc[12];
    cta::wsync();
    if (t == 1) cta::nb_arrive(1, 128);        // s' of super-stages 0, 1 | 7, 6 is in shared memory
    if (t == 3 - UP) cta::nb_arrive(2, 128);   // super-stages 2, 3 | 5
    cta::lds12(bv + SR * kc, pv);
    double val;
    if (UP) {  // the upper half produces odd stages: their s is still split into A + CC
      const double c0 = bs[SR * ko + (sp ? R_A : R_SP)];
      double c1 = 0.0;
      if (sp) c1 = bs[SR * ko + R_CC];
      val = c0 + c1;
    } else {
      val = bs[SR * ko + R_SP];
    }
    cta::tmem_get12(tq[t & 1], tc);
    if (t + 1 < NST) cta::tmem_ld12(tmr + 24 * (t + 1), tq[(t + 1) & 1]);
    cta::sched_fence();
    const double p = sub_dot12(val, tc, pv);
    if (wr && (!sp || t + 1 < NST)) bs[SR * ko + R_SP] = p;
  }
  cta::wsync();
  cta::nb_sync(5, 64);  // both halves have stored their side of s''_4
  // ---- middle: X_8 (E0) | X_9 (E1) = Sigma_4[row, 0:12] s''_8 (lanes 0..11) + Sigma_4[row, 12:24] s''_9 (lanes 16..27)
  {
    double sg[12], pv[12];
    cta::lds12(Sig4 + LS * (12 * UP + i) + 12 * sp, sg);
    cta::lds12(rec + SR * (8 + sp) + R_SP, pv);
    double d = dot12(sg, pv);
    d += cta::shfl_xor(d, 16);
    if (wr && !sp) rec[SR * (8 + UP) + R_XT + i] = d;
  }
  cta::wsync();
  cta::nb_arrive(UP ? 6 : 4, 64);
}

// back-substitution of one half: step u reads X of stage src = 8 - 2u | 9 + 2u; lanes 0..11 produce X of stage src - 2 dk (the
// next step's input), lanes 16..27 of stage src - dk
template <int UP>
QRW_HDN void state_back(double* rec, int lane, unsigned tmr) {
  constexpr int DK = 1 - 2 * UP, NST = 4 - UP;
  const int i0 = lane & 15;
  int sp = lane >> 4;
  cta::opaque(sp);
  const int i = i0 < 12 ? i0 : 11;
  const bool wr = i0 < 12;
  double* bo = rec + i - (2 - sp) * DK * SR;  // the stage this lane produces, relative to the broadcast stage
  cta::TRow tq[2];
  cta::tmem_ld12(tmr, tq[0]);
  cta::nb_sync(UP ? 6 : 4, 64);  // the elimination warp of this half has stored X_8 | X_9
#pragma unroll
  for (int u = 0; u < NST; ++u) {
    const int ks = 8 + UP - 2 * u * DK;
    double xv[12], tr[12];
    if (u > 0) cta::wsync();
    cta::lds12(rec + SR * ks + R_XT, xv);
    cta::tmem_get12(tq[u & 1], tr);
    if (u + 1 < NST) cta::tmem_ld12(tmr + 24 * (u + 1), tq[(u + 1) & 1]);
    cta::sched_fence();
    const double d = dot12(tr, xv);
    if (u == 0) cta::nb_sync(3, 128);  // the helpers' v: waited for behind the first product, not in front of it
    const double X = bo[SR * ks + R_A] - d;
    if (wr) bo[SR * ks + R_XT] = X;
  }
}

// helpers: v of the first (which = 0) / second (which = 1) stage of super-stages 0..3 (lanes 0..11) | 7..5 (lanes 16..27),
// parked in the A slot of the stage
QRW_HDN void state_helper(double* rec, int lane, unsigned tmr, int which) {
  const int i0 = lane & 15;
  int upi = lane >> 4;
  cta::opaque(upi);
  const int i = i0 < 12 ? i0 : 11;
  const bool wr = i0 < 12;
  const int k0 = (NS - 1) * upi, dk = 1 - 2 * upi;
  cta::TRow ta, tb;
  cta::tmem_ld12(tmr, ta);
  cta::tmem_ld12(tmr + 24, tb);
  cta::nb_sync(1, 128);
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int kc = k0 + 2 * t * dk, kn = kc + dk;          // the two stages of the super-stage
    const int kf = upi ? kn : kc, ksn = upi ? kc : kn;       // first and second stage (ascending)
    double m0[12], m1[12], p0[12], p1[12];
    if (t == 2) cta::nb_sync(2, 128);
    cta::lds12(rec + SR * kf + R_SP, p0);
    cta::lds12(rec + SR * ksn + R_SP, p1);
    cta::tmem_get12(ta, m0);
    cta::tmem_get12(tb, m1);
    if (t < 3) {
      cta::tmem_ld12(tmr + 48 * (t + 1), ta);
      cta::tmem_ld12(tmr + 48 * (t + 1) + 24, tb);
    }
    const double v = dot12(m0, p0) + dot12(m1, p1);
    if (wr && !(upi && t == 3)) rec[SR * (which ? ksn : kf) + R_A + i] = v;
  }
  cta::wsync();
  cta::nb_arrive(3, 128);
}

// ------------------------------------------------------------------------------------------
// One full MPC::run for one instance; every thread of the CTA calls it with the same arguments.
// Same I/O contract as mpc_solve_instance (mpc_core.cuh).  recw = the first of the FOUR warps that run the state chain
// (elimination, two helpers, back-substitution on warps recw .. recw + 3: the four lane quadrants of tensor
// memory); tmem = base address of the CTA's tensor-memory allocation (cta::TMEM_COLS columns).
// ------------------------------------------------------------------------------------------
QRW_HD void mpc_v4_solve(const MpcConst& C, double* slab, const double* xref, const double* fs, double* st_d,
                         int* st_i, int first, double* out, int* info, double* rho_out, double* dbg, int recw, unsigned tmem) {
  const Sm S(slab);
  const unsigned tm = cta::tmem_addr(tmem, 0);  // this thread's lane quadrant
  const int tid = cta::tid();
  const int k = tid >> 4, j = tid & 15, w = tid >> 5, l = tid & 31;
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
  int n_factor = 0;
  bool refactor = true;

  // ---- ADMM (osqp_solve)
  int status = ST_UNSOLVED;
  int iter = 0;
  int until_check = C.check_every;  // countdown instead of iter % check_every (an integer division per iteration)
  for (iter = 1; iter <= C.max_iter; ++iter) {
    if (refactor) {
      factor(C, S, R, Rh, S.red[120], tid, recw, tm);
      ++n_factor;
      refactor = false;
    }
    const double rhoA = eq4 ? Rh.rho_eq : Rh.rho, rinvA = eq4 ? Rh.rinv_eq : Rh.rinv;
    // (1) row values u = E (rho z - y); pair sums of the cone rows for the axis-z lane of the foot
    double u1 = 0.0, u2 = 0.0, uA = 0.0, uB = 0.0;
    if (isv) {
      u1 = R.e1 * (Rh.rho_eq * R.z1 - R.y1);
      if (bit) u2 = Ek[12 + j] * (Rh.rho_eq * 0.0 - R.y2);
      uA = R.eA * (rhoA * R.zA - R.yA);
      uB = R.eB * (Rh.rho * R.zB - R.yB);
      BAk[j] = u1;
      BDk[j] = uA + uB;
    }
    cta::wsync();
    // row j of F_k^-1 from the thread's tensor-memory lane: requested here, needed in phase (3)
    cta::TRow tfi;
    cta::tmem_ld12(tm + TM_FI + 24 * (w >> 2), tfi);
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
      const double ak = C.sigma * R.xX - R.dX * mk;
      rec[R_A + j] = ak;
      if (k == NS - 1) rec[R_SP + j] = ak;  // s of the last stage (its CC is zero)
    }
    cta::wsync();
    // (4) CC_{k-1} = D_X^{k-1} A' m_k (record of the previous stage).  An odd stage shares its warp with the stage before it: it
    // also leaves s_{k-1} = A_{k-1} + CC_{k-1} in the R_SP slot, so that the chain finds s of every EVEN stage ready
    if (isv && k > 0) {
      double v = mk;
      if (j >= 6) v += C.dt * BAk[j - 6];
      const double cc = Dk[-SR + 12 + j] * v;
      rec[-SR + R_CC + j] = cc;
      if (k & 1) rec[-SR + R_SP + j] = rec[-SR + R_A + j] + cc;
    }
    cta::bar();
    // The chain is a real call (noinline).  ptxas gives a callee the registers that are dead at the call site; with the
    // solver state of the calling thread live across it that left ~60, and the chain serialised its six broadcast loads
    // behind the FMAs for want of 12 more (SASS of r02c: one LDS.128 latency per pair of products, 170-220 cycles per
    // step against ~100).  The three chain warps therefore park their iterates in shared memory around the call.
    const int role = (w - recw) & 7;
    if (role < 6) {
      double* pk = S.park + park_off(role) + l * PK;
      pk[0] = R.xf, pk[1] = R.xX, pk[2] = R.dF, pk[3] = R.dX, pk[4] = R.z1, pk[5] = R.y1, pk[6] = R.e1, pk[7] = R.y2;
      pk[8] = R.zA, pk[9] = R.yA, pk[10] = R.eA, pk[11] = R.zB, pk[12] = R.yB, pk[13] = R.eB, pk[14] = uf, pk[15] = wv;
      const unsigned tr2 = tm + TM_ROLE + (role >= 4 ? 96 : 0);
      if (role == 0) state_elim<0>(S.SD + SB * 4, S.TT + 144, S.rec, l, tr2);
      else if (role == 4) state_elim<1>(S.SD + SB * 4, S.TT + 288 * 6, S.rec, l, tr2);
      else if (role == 1) state_back<0>(S.rec, l, tr2);
      else if (role == 5) state_back<1>(S.rec, l, tr2);
      else state_helper(S.rec, l, tm + TM_ROLE, role - 2);
      R.xf = pk[0], R.xX = pk[1], R.dF = pk[2], R.dX = pk[3], R.z1 = pk[4], R.y1 = pk[5], R.e1 = pk[6], R.y2 = pk[7];
      R.zA = pk[8], R.yA = pk[9], R.eA = pk[10], R.zB = pk[11], R.yB = pk[12], R.eB = pk[13], uf = pk[14], wv = pk[15];
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
        const double ztc = R.eA * rowa;
        const double zrc = C.alpha * ztc + (1.0 - C.alpha) * R.zA;
        double znc = zrc + rinvA * R.yA;
        if (c3 == 2) znc = dmax(znc, -25.0 * R.eA);
        znc = dmin(znc, 0.0);
        R.yA += rhoA * (zrc - znc);
        R.zA = znc;
      }
      if (c3 < 2) {
        const double ztc = R.eB * (-dxf - C.mu * fz);
        const double zrc = C.alpha * ztc + (1.0 - C.alpha) * R.zB;
        const double znc = dmin(zrc + Rh.rinv * R.yB, 0.0);
        R.yB += Rh.rho * (zrc - znc);
        R.zB = znc;
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
        yA = R.eA * R.yA;
        yB = R.eB * R.yB;
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
        row_acc(R.eA * ((c3 < 2) ? dxf - C.mu * fz : -fz), R.zA, R.eA);
        if (c3 < 2) row_acc(R.eB * (-dxf - C.mu * fz), R.zB, R.eB);
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
    st_d[ST_Z + grow(k, rowA)] = bad ? 0.0 : R.zA;
    st_d[ST_Y + grow(k, rowA)] = bad ? 0.0 : R.yA;
    if (hasB) {
      st_d[ST_Z + grow(k, rowA + 1)] = bad ? 0.0 : R.zB;
      st_d[ST_Y + grow(k, rowA + 1)] = bad ? 0.0 : R.yB;
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

}  // namespace v4
}  // namespace qrw
