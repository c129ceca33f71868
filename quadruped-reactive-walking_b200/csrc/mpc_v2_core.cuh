// mpc_v2_core.cuh -- one Solo12 centroidal-MPC solve by ONE 256-thread CTA, "force-first" KKT elimination.
//
// Same contract, same OSQP iterate sequence and same persistent state as mpc_core.cuh / mpc_cta_core.cuh; it
// replaces the same reference code:
//   MPC::run  /root/reference/src/MPC.cpp:626-649  (construct_gait :686, construct_S :665, update_ML :418,
//   update_NK :469, call_solver :501 -> OSQP, retrieve_result :569).
//
// Why a third kernel.  ncu on the v1 CTA kernel (profiles/r01_ncu_mpc_v1_cta_per_instance.txt): 128 registers
// with 534 local-memory loads per ADMM iteration, 15.2 k warp instructions per iteration, issue slots 21 % busy.
// Its block-tridiagonal solve applies a dense 24x24 pivot inverse three times per stage and iteration.  Here
//   * the reduced KKT matrix K = P + sigma I + A' diag(rho) A is eliminated FORCES FIRST: the 16 force
//     blocks F_k (12x12, independent of one another) go first; the Schur complement on the states is block
//     tridiagonal with 12x12 blocks and is factored TWISTED (stages 0..7 downwards, 15..9 upwards, stage 8
//     last), so that a solve runs two 7-step elimination chains and two 8-step back-substitution chains at
//     once in the two half-warps of one warp (17 sequential 12x12 products instead of 31).  GF = E B0 D_f F^-1
//     and M = GF D_f B0' E are formed at factorisation time: one broadcast of the force right-hand side gives
//     u = F^-1 t and what the dynamics rows see of it, one broadcast of the six dynamics-row residuals gives
//     x~_f and z~.  About 5 k warp instructions per ADMM iteration instead of 15 k;
//   * 16 lanes per horizon stage: lane j < 12 of stage k owns force f_k[j], state X_{k+1}[j], dynamics row
//     (k,j), force-enable row (k,j) and up to two friction-cone rows of its foot; lanes 12..15 idle.  The
//     iterates live in registers of their owner; everything else a stage owns (exchange buffers, D, E, B_k,
//     F^-1, GF, M) sits in one shared-memory record.  Stage-internal exchanges use __syncwarp; only the
//     state chain (a noinline call, one warp) needs the two CTA barriers of an iteration.
//   * 89 KB of shared memory and 128 registers: two CTAs (16 warps) per SM.
//   * (round 2) the matrix rows a lane reads again in every iteration live in TENSOR MEMORY, not in shared memory:
//     F_k^-1 (row j of its owner thread) and the rows / columns of T_k, U_k, Sig_k the three chain warps walk through.
//     ncu on round 1 (profiles/r01f_ncu_mpc_v2.txt) had the SM's shared-memory pipe at 0.51 wavefronts per cycle and the
//     chain -- a sequence of dependent shared-memory round trips -- 1.8x slower inside the kernel than alone, because the
//     co-resident CTA's LDS bursts queue in front of it.  tcgen05.ld (32x32b: a thread reads columns of its own TMEM
//     lane) goes through a separate pipe: measured 840 B/cycle/SM, no interference with LDS in either direction
//     (tools/tmem_bench.cu, profiles/r02_tmem_bench.txt).  256 columns per CTA (two CTAs share the SM's 512): columns
//     [0, 48) hold the F^-1 rows of the two warps of a lane quadrant, [48, 240) the eight rows of the chain role that
//     runs in that quadrant.  No tcgen05.mma anywhere: the tensor cores have no f64 kind, their memory is what is used.
// It is still an exact solve of the same KKT system: the iterates differ from OSQP's by round-off only.
#pragma once
#include "mpc_common.cuh"

namespace qrw {
namespace v2 {

#ifdef QRW_V2_TIMING
// clock read that the compiler may not move across barriers or shared-memory traffic (probe build only)
__device__ __forceinline__ long long qrw_clock() {
  long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory");
  return t;
}
#endif

constexpr int NT = 256;  // threads per CTA
constexpr int LD = 14;   // padded row length of the 12x12 matrices in shared memory (conflict-free LDS.128)
constexpr int MS = 12 * LD;

// shared slab layout (doubles).  Everything a stage owns sits in one RECORD of SR doubles, so that a thread
// addresses all of it (and the records of the neighbouring stages) from one base register.
constexpr int V_SIG = 0;              // [16][12][14] Sig_k  (during factorisation: G_kk)
constexpr int V_T = V_SIG + 16 * MS;  // [15][12][14] T_k    (during factorisation: G_{k,k+1})
constexpr int V_REC = V_T + 15 * MS;  // [16][SR]     stage records
constexpr int R_BA = 0, R_BB = 16, R_BC = 32, R_BD = 48;  // 4 x 16 stage-internal exchange buffers
constexpr int R_A = 64;    // [12] sigma x_X - D_X m_k            (written by the stage's own lanes)
constexpr int R_CC = 76;   // [12] D_X A' m_{k+1}                 (written by the lanes of stage k+1)
constexpr int R_XT = 88;   // [12] x~_X of the stage
constexpr int R_SP = 100;  // [12] forward-recurrence result s'_k
constexpr int R_D = 112;   // [24] Ruiz column scaling: 0..11 force, 12..23 state
constexpr int R_E = 136;   // [44] Ruiz row scaling: 12 dyn | 12 enable | 20 cone
constexpr int R_BANG = 180;  // [3][12] angular block of B_k (unscaled)
constexpr int R_B1 = 216;  // [12] scaled right-hand side of the dynamics rows
constexpr int R_GF = 228;  // [6][14] rows 6..11 of GF = E B0 D_f F^-1 (what the dynamics rows see of a force right-hand side)
constexpr int R_MR = 312;  // [6][6]  M = GF D_f B0' E (rows 6..11 x rows 6..11)
constexpr int SR = 348;    // (F_k^-1 left the record for tensor memory in round 2)
// tensor-memory columns of a lane quadrant (32-bit columns; a row of 12 doubles = 24 columns)
constexpr int TM_FI = 0;     // [2][24] row j of F_k^-1 of the thread: warps 0..3 at column 0, warps 4..7 at column 24
constexpr int TM_ROLE = 48;  // [8][24] rows of the chain role of this quadrant's warp (elimination / helper / back-substitution)
static_assert(TM_ROLE + 8 * 24 <= cta::TMEM_COLS, "tensor-memory columns of a CTA");
constexpr int V_W = V_REC + 16 * SR;  // [144] Gauss-Jordan ping-pong / T_k staging of the factorisation
constexpr int V_RED = V_W + 144;      // 128   reductions
constexpr int V_INT = V_RED + 128;    // 128 ints
constexpr int V_PARK = V_INT + 64;    // [3][32][PK] solver state of the three chain warps while they run the chain
constexpr int PK = 18;                // 16 doubles per lane, padded so that the lanes' 16-byte accesses fall on distinct banks
constexpr int V2_SLAB_DOUBLES = V_PARK + 3 * 32 * PK;
constexpr int V2_SLAB_BYTES = V2_SLAB_DOUBLES * 8;
// factorisation scratch inside a record, aliased on the iteration-only vectors [0, 112)
constexpr int R_Y = 0;   // [12][6] F_k^-1 (D_f B0')(:, rows 6..11); dead after step (c)
constexpr int R_N = 72;  // [36]    N_k (6x6); dead after step (d)
static_assert(R_N + 36 <= R_D, "factorisation scratch overruns the record");
static_assert((SR * 8) % 16 == 0 && (V_REC * 8) % 16 == 0, "records must keep 16-byte alignment");
static_assert(V2_SLAB_BYTES <= 112 * 1024, "two CTAs per SM need <= 113 KB each");
constexpr int V2_TMEM_SLOT = 120;  // index in the int buffer where tcgen05.alloc parks the CTA's tensor-memory base address

struct Sm {
  double *Sig, *T, *rec, *W, *red, *park;
  int* ib;
  QRW_HD explicit Sm(double* b)
      : Sig(b + V_SIG), T(b + V_T), rec(b + V_REC), W(b + V_W), red(b + V_RED), park(b + V_PARK),
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
//   G_kk, G_{k,k+1}: Schur complement on the states       -> Sig / T slots, then the twisted block LDL'
//   (pivot inverses Sig_k, coupling matrices T_k = Sig_k G_{k,k+1} for k <= 7, U_k = Sig_k G_{k-1,k}' for k >= 9)
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
  double* W = S.W;
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
  // ---- (d) rows of G_kk (into the Sig slot) and G_{k,k+1} (into the T slot)
  if (isv) {
    const int i = j;
    const double* En = Ek + SR;  // dynamics-row scaling of stage k+1 (valid for k < 15)
    const double* Dn = Dk + SR;  // column scaling of stage k+1
    const double* Nn = N + SR;
    const double r2 = Rh.rho_eq * Rh.rho_eq;
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
      S.Sig[MS * k + LD * i + ip] = v;
      if (k < NS - 1) S.T[MS * k + LD * i + ip] = o;
    }
  }
  cta::bar();
  // ---- (e) TWISTED block LDL' of the state system (144 threads, one matrix entry per thread): stages 15..9 are
  // eliminated upwards (U_s = Sig_s G_{s-1,s}' into T slot s-1), stages 0..7 downwards (T_s = Sig_s G_{s,s+1} into
  // T slot s), stage 8 last with both Schur updates.  The solve then runs two chains of 7 steps at once.
  const int ei = tid / 12, ej = tid - 12 * ei;  // entry (ei, ej) for tid < 144
  const bool ent = tid < 144;
  for (int step = 0; step < NS; ++step) {
    // order: 15, 14, ..., 9 (bottom), 0, 1, ..., 7 (top), 8 (middle)
    const int sg = step < 7 ? NS - 1 - step : (step < 15 ? step - 7 : 8);
    const bool bottom = step < 7;
    double* G = S.Sig + MS * sg;
    // Gauss-Jordan inverse, ping-pong G <-> W (12 pivots: the result is back in G)
    for (int p = 0; p < 12; ++p) {
      const double* Wc = (p & 1) ? W : G;
      double* Wn = (p & 1) ? G : W;
      const int ldc = (p & 1) ? 12 : LD, ldn = (p & 1) ? LD : 12;
      if (ent) {
        const double d = 1.0 / Wc[ldc * p + p];
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
    if (sg == 8) break;
    const int so = bottom ? sg - 1 : sg;          // T slot holding the coupling block G_{so,so+1}
    const int st = bottom ? sg - 1 : sg + 1;      // stage whose pivot block receives the Schur update
    double* O = S.T + MS * so;
    // coupling matrix Sig_s G' (bottom) or Sig_s G (top), staged in W
    if (ent) {
      double acc = 0.0;
#pragma unroll
      for (int m = 0; m < 12; ++m) acc += G[LD * ei + m] * (bottom ? O[LD * ej + m] : O[LD * m + ej]);
      W[12 * ei + ej] = acc;
    }
    cta::bar();
    double tv = 0.0;
    if (ent) {
      double acc = 0.0;
#pragma unroll
      for (int m = 0; m < 12; ++m) acc += (bottom ? O[LD * ei + m] : O[LD * m + ei]) * W[12 * m + ej];
      S.Sig[MS * st + LD * ei + ej] -= acc;
      tv = W[12 * ei + ej];
    }
    cta::bar();
    if (ent) O[LD * ei + ej] = tv;
  }
  // the scratch overlapped CC of the last stage, which nobody writes and the recurrence reads as zero
  if (tid < 12) S.rec[SR * (NS - 1) + R_CC + tid] = 0.0;
  cta::bar();
  // ---- (f) the three chain warps copy the rows they walk through in every iteration into their tensor-memory lanes
  // (TM_ROLE + 24 t = what step t of the role reads); lanes 12..15 of a half-warp duplicate row 11
  const int w = tid >> 5, role = (w - recw) & 7;
  if (role < 3) {
    const int lane = tid & 31, upi = lane >> 4, i0 = lane & 15, i = i0 < 12 ? i0 : 11;
    const int k0 = (NS - 1) * upi, dk = 1 - 2 * upi;
    for (int t = 0; t < 8; ++t) {
      double c[12];
      if (role == 0) {
        // elimination: column i of T_t | U_{15-t} (t < 7), of T_7 for the middle step
        const int slot = t < 7 ? k0 + dk * t - upi : 7;
        cta::lds12c<LD>(S.T + MS * slot + i, c);
      } else if (role == 1) {
        // helper: row i of Sig_t | Sig_{15-t}, t < 7
        cta::lds12(S.Sig + MS * (k0 + dk * (t < 7 ? t : 6)) + LD * i, c);
      } else {
        // back-substitution: row i of T_7, .., T_0 | U_9, .., U_15 (the upper half repeats its last row)
        const int k2 = t == 0 ? 7 + 2 * upi : (t < 7 ? 7 + 2 * upi - dk * t : k0);
        cta::lds12(S.T + MS * (k2 - upi) + LD * i, c);
      }
      cta::tmem_st12(tm + TM_ROLE + 24 * t, c);
    }
  }
  cta::tmem_wait_st();
}

// ------------------------------------------------------------------------------------------
// State system of one KKT solve with the twisted factorisation, ONE warp, fully unrolled.  Lanes 0..11 run the
// chain of stages 0..7 (downwards), lanes 16..27 the chain of stages 15..8 (upwards), in lockstep:
//   s_k = A_k + CC_k
//   elimination, 7 steps : s'_{k+1} = s_{k+1} - T_k' s'_k  |  s'_{k-1} = s_{k-1} - U_k' s'_k ;  v_k = Sig_k s'_k
//   middle               : s''_8 = s'_8 - T_7' s'_7 ,  v_7 = Sig_7 s'_7 ,  X_8 = Sig_8 s''_8
//   back-substitution    : X_k = v_k - T_k X_{k+1} (k = 7..0)  |  X_k = v_k - U_k X_{k-1} (k = 9..15)
// T_k sits in T slot k (k <= 7), U_k in T slot k-1 (k >= 9).  The vector of a step is broadcast through the stage
// record (STS, __syncwarp, LDS.128); v_k is parked in the A slot of its stage.
// ------------------------------------------------------------------------------------------
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
// Scheduling: a step is one dependent sequence  STS -> LDS.128 x6 (the broadcast) -> 3 DFMA deep -> 3 DADD -> STS.
// The matrix column / row of step t+1 comes from the lane's tensor memory: it is requested in step t, right behind the
// wait for the row of step t (tcgen05.wait::ld waits for everything outstanding, so one request is in flight at a time),
// and arrives under the FMAs without touching the shared-memory queue the broadcast of the next step goes through.
// Three warps share the work (round 2; each reads its rows from its own lane quadrant of tensor memory):
//   E (warp recw)     elimination and the middle step; hands s'_k to H (named barriers 1, 2) and X_8 | v_7 to S (barrier 4)
//   H (warp recw + 1) v_k = Sig_k s'_k for k = 0..6 | 15..9, behind the elimination (barrier 3 hands them to S)
//   S (warp recw + 2) back-substitution
#ifdef QRW_V2_TIMING
// clock read that cannot be issued before x is known (predicated on a comparison of x that is practically always true)
__device__ __forceinline__ long long qrw_clock_after(double x) {
  long long t = 0;
  asm volatile("{ .reg .pred q; setp.ne.f64 q, %1, 0d7E37E43C8800759C; @q mov.u64 %0, %%clock64; }" : "+l"(t) : "d"(x) : "memory");
  return t;
}
__shared__ long long chain_stamp[8];  // probe build: clock at the five seams of the chain (lane 0 of E and of S)
#define QRW_STAMP(n, x) { const long long t_ = qrw_clock_after(x); if (lane == 0) chain_stamp[n] = t_; }
#else
#define QRW_STAMP(n, x) ((void)0)
#endif
QRW_HDN void state_elim(const double* Sig, double* rec, int lane, unsigned tmr) {
  double lastv = (double)lane;  // what the probe build hangs its clock reads on
  QRW_STAMP(0, lastv);
  const int i0 = lane & 15;
  int upi = lane >> 4;  // 0: lower half of the warp (stages 0..7), 1: upper half (stages 15..8)
  cta::opaque(upi);     // keeps ptxas from cloning the whole chain per half (it did: 468 DFMA instead of 288, run one after the other)
  const bool up = upi != 0;
  const int i = i0 < 12 ? i0 : 11;
  const bool wr = i0 < 12;
  // stage of this half at elimination step t: lower t, upper 15 - t
  const int k0 = (NS - 1) * upi;
  const int dk = 1 - 2 * upi;
  cta::TRow tq[2];  // the column in flight; two buffers so that the request does not wait for the registers of the column in use
  double sn;
  cta::tmem_ld12(tmr, tq[0]);  // column i of T_0 / U_15
  {
    double* r0 = rec + SR * k0;
    double* r1 = rec + SR * (k0 + dk);
    const double p = r0[R_A + i] + r0[R_CC + i];
    if (wr) r0[R_SP + i] = p;
    sn = r1[R_A + i] + r1[R_CC + i];
  }
  double tc[12];
#pragma unroll
  for (int t = 0; t < 7; ++t) {
    const int kc = k0 + dk * t;  // stage whose s' is broadcast
    const int kn = kc + dk;      // stage produced
    double pv[12], sn2 = 0.0;
    cta::wsync();
    if (t == 3) cta::nb_arrive(1, 64);  // s'_0..3 | s'_15..12 are in shared memory: the helper warp starts on their v_k
    if (t == 6) cta::nb_arrive(2, 64);  // s'_4..6 | s'_11..9
    cta::lds12(rec + SR * kc + R_SP, pv);
    cta::tmem_get12(tq[t & 1], tc);
    cta::tmem_ld12(tmr + 24 * (t + 1), tq[(t + 1) & 1]);  // column i of T_k / U_k of the next step (t = 6: of T_7 on every lane)
    if (t < 6) {
      const double* r2 = rec + SR * (kn + dk);
      sn2 = r2[R_A + i] + r2[R_CC + i];
    }
    cta::sched_fence();
    const double p = sub_dot12(sn, tc, pv);
    if (wr) rec[SR * kn + R_SP + i] = p;
    lastv = p;
    sn = sn2;
  }
  QRW_STAMP(1, lastv);
  // ---- middle: s''_8 = s'_8 - T_7' s'_7 on every lane, stored by the lower half
  {
    double* r8 = rec + SR * 8;
    double pv[12];
    cta::wsync();
    cta::lds12(rec + SR * 7 + R_SP, pv);
    const double s8p = r8[R_SP + i];
    cta::tmem_get12(tq[1], tc);
    cta::sched_fence();
    const double s8 = sub_dot12(s8p, tc, pv);
    cta::wsync();  // everybody has read s'_8 before it is replaced by s''_8
    if (wr && !up) r8[R_SP + i] = s8;
    lastv = s8;
    cta::wsync();
  }
  QRW_STAMP(2, lastv);
  // ---- v_7 = Sig_7 s'_7 (lower half) and X_8 = Sig_8 s''_8 (upper half); the other v_k come from the helper warp
  {
    double sg[12], pv[12];
    const int kc = 7 + upi;
    cta::lds12(Sig + MS * kc + LD * i, sg);
    cta::lds12(rec + SR * kc + R_SP, pv);
    const double v = dot12(sg, pv);
    if (wr) rec[SR * kc + R_A + (R_XT - R_A) * upi + i] = v;
    lastv = v;
  }
  cta::wsync();
  cta::nb_arrive(4, 64);  // X_8 and v_7 are in shared memory: the back-substitution warp starts
  QRW_STAMP(3, lastv);
}

// back-substitution: lower half stages 7..0, upper half stages 9..15 (idle in the last step):  X_k = v_k - T_k X_{k+1}
QRW_HDN void state_back(double* rec, int lane, unsigned tmr) {
  const int i0 = lane & 15;
  int upi = lane >> 4;
  cta::opaque(upi);
  const bool up = upi != 0;
  const int i = i0 < 12 ? i0 : 11;
  const bool wr = i0 < 12;
  const int k0 = (NS - 1) * upi;
  const int dk = 1 - 2 * upi;
  cta::TRow tq[2];
  double tr[12], vk = 0.0, lastv = 0.0;
  cta::tmem_ld12(tmr, tq[0]);  // row i of T_7 / U_9
  cta::nb_sync(4, 64);         // the elimination warp has stored X_8 and v_7
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int kk = u < 7 ? 7 + 2 * upi - dk * u : k0;  // stage produced: 7 - u | 9 + u, the upper half repeats 15
    const int kp = kk + dk;                            // stage whose X is broadcast
    double xv[12], vn = 0.0;
    if (u > 0) cta::wsync();
    cta::lds12(rec + SR * kp + R_XT, xv);
    cta::tmem_get12(tq[u & 1], tr);
    if (u < 7) {
      const int k2 = u < 6 ? 7 + 2 * upi - dk * (u + 1) : k0;
      cta::tmem_ld12(tmr + 24 * (u + 1), tq[(u + 1) & 1]);  // row i of T_k / U_k of the next step
      if (u > 0) vn = rec[SR * k2 + R_A + i];
    }
    cta::sched_fence();
    double X;
    if (u == 0) {
      // the first step waits for the helper's v_k behind its product, not in front of it
      const double d = dot12(tr, xv);
      cta::nb_sync(3, 64);
      vk = rec[SR * kk + R_A + i];
      vn = rec[SR * (7 + 2 * upi - dk) + R_A + i];
      X = vk - d;
    } else {
      X = sub_dot12(vk, tr, xv);
    }
    if (wr && (!up || u < 7)) rec[SR * kk + R_XT + i] = X;
    lastv = X;
    vk = vn;
  }
  QRW_STAMP(4, lastv);
}

// The helper warp: v_k = Sig_k s'_k for k = 0..6 (lower half) and 15..9 (upper half), parked in the A slot of the stage, in
// two groups behind the elimination (named barriers 1 and 2); barrier 3 hands them to the back-substitution.  Seven
// independent products; the Sig rows come from tensor memory, one request in flight.
QRW_HDN void state_helper(double* rec, int lane, unsigned tmr) {
  const int i0 = lane & 15;
  int upi = lane >> 4;
  cta::opaque(upi);
  const int i = i0 < 12 ? i0 : 11;
  const bool wr = i0 < 12;
  const int k0 = (NS - 1) * upi;
  const int dk = 1 - 2 * upi;
  cta::TRow tq[2];
  double sg[12], pv[2][12];
  cta::tmem_ld12(tmr, tq[0]);
  cta::nb_sync(1, 64);
  cta::lds12(rec + SR * k0 + R_SP, pv[0]);
#pragma unroll
  for (int t = 0; t < 7; ++t) {
    const int kc = k0 + dk * t;
    cta::tmem_get12(tq[t & 1], sg);
    if (t < 6) {
      const int k2 = kc + dk;
      cta::tmem_ld12(tmr + 24 * (t + 1), tq[(t + 1) & 1]);
      if (t == 3) cta::nb_sync(2, 64);
      cta::lds12(rec + SR * k2 + R_SP, pv[(t + 1) & 1]);
    }
    const double v = dot12(sg, pv[t & 1]);
    if (wr) rec[SR * kc + R_A + i] = v;
  }
  cta::wsync();
  cta::nb_arrive(3, 64);
}

// ------------------------------------------------------------------------------------------
// One full MPC::run for one instance; every thread of the CTA calls it with the same arguments.
// Same I/O contract as mpc_solve_instance (mpc_core.cuh).  recw = the first of the three warps that run the state chain
// (elimination, helper, back-substitution on warps recw, recw + 1, recw + 2: three different lane quadrants of tensor
// memory); tmem = base address of the CTA's tensor-memory allocation (cta::TMEM_COLS columns).
// ------------------------------------------------------------------------------------------
QRW_HD void mpc_v2_solve(const MpcConst& C, double* slab, const double* xref, const double* fs, double* st_d,
                         int* st_i, int first, double* out, int* info, double* rho_out, double* dbg, int recw, unsigned tmem) {
  const Sm S(slab);
  const unsigned tm = cta::tmem_addr(tmem, 0);  // this thread's lane quadrant
#ifdef QRW_V2_TIMING
  const long long t_entry = qrw_clock();
#endif
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

#ifdef QRW_V2_TIMING
  // probe build (tools/timing_probe.py, -DQRW_V2_TIMING): cycles of the segments of an iteration seen by thread 0 of a warp
  // that does NOT run the chain (0: phases 1-4, 3: phases 5-7), by lane 0 of the chain warp (4: the chain), and per solve
  // (5: everything before the loop, 6: factorisations, 7: residual checks).  Parked in the last entries of the debug dump.
  long long tq[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long tp[7] = {0, 0, 0, 0, 0, 0, 0};  // phases (1)..(7) one by one
  long long tcs[4] = {0, 0, 0, 0};          // inside the chain: elimination, middle, v products, back-substitution
  const int tw = ((recw + 1) & 7) * 32;
  tq[5] = qrw_clock() - t_entry;
#define QRW_TICK(i, t0) tq[i] += qrw_clock() - (t0)
#else
#define QRW_TICK(i, t0) ((void)0)
#endif
  // ---- ADMM (osqp_solve)
  int status = ST_UNSOLVED;
  int iter = 0;
  int until_check = C.check_every;  // countdown instead of iter % check_every (an integer division per iteration)
  for (iter = 1; iter <= C.max_iter; ++iter) {
#ifdef QRW_V2_TIMING
    long long tA = qrw_clock();
    long long tL = tA;
#endif
    if (refactor) {
#ifdef QRW_V2_TIMING
      const long long tF = qrw_clock();
#endif
      factor(C, S, R, Rh, S.red[120], tid, recw, tm);
      QRW_TICK(6, tF);
#ifdef QRW_V2_TIMING
      tA = qrw_clock();
      tL = tA;
#endif
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
#ifdef QRW_V2_TIMING
    tp[0] += qrw_clock() - tL, tL = qrw_clock();
#endif
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
#ifdef QRW_V2_TIMING
    tp[1] += qrw_clock() - tL, tL = qrw_clock();
#endif
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
      rec[R_A + j] = C.sigma * R.xX - R.dX * mk;
    }
    cta::wsync();
#ifdef QRW_V2_TIMING
    tp[2] += qrw_clock() - tL, tL = qrw_clock();
#endif
    // (4) CC_{k-1} = D_X^{k-1} A' m_k (record of the previous stage)
    if (isv && k > 0) {
      double v = mk;
      if (j >= 6) v += C.dt * BAk[j - 6];
      rec[-SR + R_CC + j] = Dk[-SR + 12 + j] * v;
    }
    QRW_TICK(0, tA);
#ifdef QRW_V2_TIMING
    tp[3] += qrw_clock() - tL;
#endif
    cta::bar();
#ifdef QRW_V2_TIMING
    tA = qrw_clock();
#endif
    // The chain is a real call (noinline).  ptxas gives a callee the registers that are dead at the call site; with the
    // solver state of the calling thread live across it that left ~60, and the chain serialised its six broadcast loads
    // behind the FMAs for want of 12 more (SASS of r02c: one LDS.128 latency per pair of products, 170-220 cycles per
    // step against ~100).  The three chain warps therefore park their iterates in shared memory around the call.
    const int role = (w - recw) & 7;
    if (role < 3) {
      double* pk = S.park + (32 * role + l) * PK;
      pk[0] = R.xf, pk[1] = R.xX, pk[2] = R.dF, pk[3] = R.dX, pk[4] = R.z1, pk[5] = R.y1, pk[6] = R.e1, pk[7] = R.y2;
      pk[8] = R.zA, pk[9] = R.yA, pk[10] = R.eA, pk[11] = R.zB, pk[12] = R.yB, pk[13] = R.eB, pk[14] = uf, pk[15] = wv;
      if (role == 0) state_elim(S.Sig, S.rec, l, tm + TM_ROLE);
      else if (role == 1) state_helper(S.rec, l, tm + TM_ROLE);
      else state_back(S.rec, l, tm + TM_ROLE);
      R.xf = pk[0], R.xX = pk[1], R.dF = pk[2], R.dX = pk[3], R.z1 = pk[4], R.y1 = pk[5], R.e1 = pk[6], R.y2 = pk[7];
      R.zA = pk[8], R.yA = pk[9], R.eA = pk[10], R.zB = pk[11], R.yB = pk[12], R.eB = pk[13], uf = pk[14], wv = pk[15];
    }
    cta::bar();
#ifdef QRW_V2_TIMING
    if (w == recw) {  // the stamps of the elimination (0..3) and of the back-substitution warp (4) are complete behind the barrier
      const volatile long long* cs = chain_stamp;
      tq[4] += cs[4] - cs[0];
      for (int q = 0; q < 4; ++q) tcs[q] += cs[q + 1] - cs[q];
    }
    tA = qrw_clock();
    tL = tA;
#endif
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
#ifdef QRW_V2_TIMING
    tp[4] += qrw_clock() - tL, tL = qrw_clock();
#endif
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
#ifdef QRW_V2_TIMING
    tp[5] += qrw_clock() - tL, tL = qrw_clock();
#endif
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
    QRW_TICK(3, tA);
#ifdef QRW_V2_TIMING
    tp[6] += qrw_clock() - tL;
#endif
    if (--until_check == 0) {
      until_check = C.check_every;
      // residuals (auxil.c compute_pri_res / compute_dua_res / compute_*_tol)
#ifdef QRW_V2_TIMING
      const long long tC = qrw_clock();
#endif
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
      // The termination test needs the six UNSCALED norms (odd q); the scaled ones (even q) feed the rho estimate only, on every
      // (adapt_every / check_every)-th check: their reductions are skipped otherwise.  Across the warps, lane q reduces column
      // q (same order as before: warp 0, 1, .. 7) and broadcasts it, instead of every thread reading all 96 partial maxima.
      const bool adapt_now = (iter % C.adapt_every == 0);
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        if (adapt_now || (q & 1)) {
          m[q] = cta::warp_max_x(m[q]);
          if (l == 0) S.red[16 + 12 * w + q] = m[q];
        }
      }
      cta::bar();
      {
        const int lq = l < 12 ? l : 11;
        double v = S.red[16 + lq];
#pragma unroll
        for (int ww = 1; ww < NT / 32; ++ww) v = dmax(v, S.red[16 + 12 * ww + lq]);
#pragma unroll
        for (int q = 0; q < 12; ++q)
          if (adapt_now || (q & 1)) m[q] = cta::bcast(v, q);
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
      if (adapt_now) {
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
      QRW_TICK(7, tC);
    }
  }
  if (iter > C.max_iter) iter = C.max_iter;
  if (status == ST_UNSOLVED) status = ST_MAX_ITER;

#ifdef QRW_V2_TIMING
  if (dbg) {
    if (tid == tw) {
      dbg[1272] = (double)tq[0], dbg[1275] = (double)tq[3];
      for (int i = 5; i < 8; ++i) dbg[1273 + i] = (double)tq[i];
      for (int i = 0; i < 7; ++i) dbg[1260 + i] = (double)tp[i];
    }
    if (w == recw && l == 0) {
      dbg[1276] = (double)tq[4];
      for (int q = 0; q < 4; ++q) dbg[1267 + q] = (double)tcs[q];
    }
    if (tid == 0) dbg[1277] = (double)(iter > C.max_iter ? C.max_iter : iter);
  }
#endif
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

}  // namespace v2
}  // namespace qrw
