// mpc_common.cuh -- definitions shared by every MPC kernel core: problem dimensions, the persistent per-robot state layout
// in HBM, OSQP status codes, the constants MPC::MPC hard-codes (/root/reference/src/MPC.cpp:17-29,330,346-348,527-539), the
// rho set of OSQP.  The shipped library compiles ONE MPC kernel (mpc_v2_core.cuh); the earlier cores (mpc_core.cuh warp per
// robot, mpc_cta_core.cuh 24x24 block CTA) and the nested-dissection experiment (mpc_v3_core.cuh) are A/B-test builds
// (-DQRW_AB_KERNELS) and include this header too.
#pragma once
#include "qrw_spmd.cuh"

namespace qrw {

constexpr int NS = 16;    // horizon steps (T_mpc / dt_mpc of src/config_solo12.yaml:11-13)
constexpr int NX = 384;   // 24 * NS variables
constexpr int NR = 704;   // 44 * NS constraint rows
constexpr int NGMAX = 32; // max rows of the gait / fsteps matrices


// per-instance persistent state in HBM (what OSQP's workspace + the MPC object carry between
// calls: scaled iterates, rho, stale B_k values, gait and S_gait)
constexpr int ST_X = 0, ST_Z = NX, ST_Y = NX + NR, ST_BANG = NX + 2 * NR, ST_RHO = ST_BANG + 576;
constexpr int MPC_STATE_DOUBLES = 2376;  // 2369 rounded up to a multiple of 8
constexpr int MPC_STATE_INTS = 64;       // gait bits [0,32) | S bits [32,48) | spare

enum MpcStatus {  // OSQP status values
  ST_SOLVED = 1,
  ST_SOLVED_INACCURATE = 2,
  ST_MAX_ITER = -2,
  ST_NON_CVX = -7,
  ST_UNSOLVED = -10
};

struct MpcConst {
  double dt, dtm, mu, g8, sigma, alpha, eps_abs, eps_rel, rho0, adapt_tol;
  double w[12];       // state weights            (MPC.cpp:330)
  double wf;          // force weight 5e-5f       (MPC.cpp:346-348)
  double gI[9];       // body inertia, row-major  (MPC.cpp:25-26)
  double foot[12];    // default footholds 3 x 4  (MPC.cpp:24)
  double off_com[3];  // (0, 0, -0.03)            (MPC.cpp:21)
  int N_gait, max_iter, check_every, adapt_every;
};


// global index of stage-local variable a of stage k: a<12 -> force f_k[a], else X_{k+1}[a-12]
QRW_HD int gvar(int k, int a) { return a < 12 ? 192 + 12 * k + a : 12 * k + (a - 12); }


struct RhoSet {
  double rho, rho_eq, rinv, rinv_eq;
  QRW_HD explicit RhoSet(double r) {
    rho = dmin(dmax(r, 1e-6), 1e6);  // RHO_MIN / RHO_MAX
    rho_eq = 1e3 * rho;              // RHO_EQ_OVER_RHO_INEQ
    rinv = 1.0 / rho;
    rinv_eq = 1.0 / rho_eq;
  }
};

}  // namespace qrw

// ------------------------------------------------------------------------------------------
// Host-side constants (the reference hard-codes them: MPC.cpp:17-29,330,346-348,527-539).
// Float literals are widened exactly as the C++ does (e.g. 0.9f -> 0.8999999761581421).
// ------------------------------------------------------------------------------------------
namespace qrw {
inline MpcConst make_mpc_const(double dt, int N_gait) {
  MpcConst C;
  const double mass = 2.50000279f;
  C.dt = dt;
  C.dtm = dt / mass;
  C.mu = 0.9f;
  C.g8 = -9.81f * dt;
  C.sigma = 1e-6;
  C.alpha = 1.6;
  C.eps_abs = 1e-6;
  C.eps_rel = 1e-6;
  C.rho0 = 0.1;
  C.adapt_tol = 5.0;
  const double w[12] = {2.0f, 2.0f, 20.0f, 0.25f, 0.25f, 10.0f, 0.2f, 0.2f, 0.2f, 0.0f, 0.0f, 0.3f};
  for (int i = 0; i < 12; ++i) C.w[i] = w[i];
  C.wf = 5e-5f;
  const double gi[9] = {3.09249e-2, -8.00101e-7, 1.865287e-5, -8.00101e-7, 5.106100e-2,
                        1.245813e-4, 1.865287e-5, 1.245813e-4, 6.939757e-2};
  for (int i = 0; i < 9; ++i) C.gI[i] = gi[i];
  const double fh[12] = {0.19, 0.19, -0.19, -0.19, 0.15005, -0.15005, 0.15005, -0.15005, 0.0, 0.0, 0.0, 0.0};
  for (int i = 0; i < 12; ++i) C.foot[i] = fh[i];
  C.off_com[0] = 0.0;
  C.off_com[1] = 0.0;
  C.off_com[2] = -0.03;
  C.N_gait = N_gait;
  C.max_iter = 4000;
  C.check_every = 25;
  C.adapt_every = 200;
  return C;
}
}  // namespace qrw
