// planner_core.cuh -- the reference's input producers for B robots:
//   Gait::updateGait                    /root/reference/src/Gait.cpp:189-260 (+ create_* :39-109, getPhaseDuration :142-187)
//   FootstepPlanner::updateFootsteps    /root/reference/src/FootstepPlanner.cpp:52-241
//   FootTrajectoryGenerator::update     /root/reference/src/FootTrajectoryGenerator.cpp:41-151
//   StatePlanner::computeReferenceStates /root/reference/src/StatePlanner.cpp:22-60
//   feet commands in the base frame     /root/reference/scripts/Controller.py:286-296
// in the order scripts/Controller.py:222-236 calls them, selected by the stage bits below, so that xref / fsteps of the
// MPC and the contacts / foot goals of the whole-body tick are produced in HBM next to their consumers.
//
// Thread mapping: FOUR lanes per robot, lane j = foot j (FL, FR, HL, HR).  A lane owns everything indexed by its foot
// (footstep columns 3j..3j+2, its swing polynomial, its target) and the xref columns j, j+4, ...; the 0/1 gait matrices
// are three 128-bit nibble arrays (row r = bits 4r..4r+3, bit j = foot j) that every lane of the robot updates
// redundantly in registers and lane 0 writes back.  No lane reads what another lane of the same tick writes; the only
// ordering constraint -- the gait words are read by all four lanes and written by lane 0 -- is met by running every
// lane's `compute` before any lane's `store` (a __syncwarp in the kernel, two passes in the CPU emulation build).
//
// Joystick codes: 1 pacing, 2 bounding, 3 trot, 4 static as Gait::changeGait (Gait.cpp:199-221); code 5 selects
// Gait::create_walk (Gait.cpp:39-55), which the reference defines but never calls -- an extension so that BASELINE
// config 3's walk schedule can be produced on the GPU.  Any other code leaves the desired gait alone.
//
// HBM layout ("field-major"): field f of length L_f per robot lives at st + B*off_f + b*L_f, so the four lanes of a
// robot and the eight robots of a warp touch one contiguous run per field.
#pragma once
#include "qrw_spmd.cuh"

#if QRW_ON_GPU
#define QRW_HDB __host__ __device__ __forceinline__
#else
#define QRW_HDB inline
#endif

namespace qrw {
namespace plan {

enum { ST_GAIT = 1, ST_FOOTSTEPS = 2, ST_FOOTTRAJ = 4, ST_STATE = 8, ST_FEETCMD = 16, ST_ALL = 31 };

// per-robot double fields (offsets in doubles, multiplied by B)
enum {
  F_CURFS = 0,     // FootstepPlanner::currentFootstep_   [4][3]
  F_TARGET = 12,   // FootstepPlanner::targetFootstep_    [4][3]
  F_OTARGET = 24,  // FootstepPlanner::o_targetFootstep_  [4][3]
  F_FTGTGT = 36,   // FootTrajectoryGenerator::targetFootstep_
  F_FPOS = 48,     // position_ / velocity_ / acceleration_
  F_FVEL = 60,
  F_FACC = 72,
  F_PCMD = 84,  // Controller.feet_p_cmd / feet_v_cmd / feet_a_cmd
  F_VCMD = 96,
  F_ACMD = 108,
  F_AX = 120,  // Ax, Ay [4][6]
  F_AY = 144,
  F_T0S = 168,  // t0s, t_swing [4]
  F_TSW = 172,
  F_CONTACTS = 176,  // currentGait_.row(0) as doubles [4] (output for the whole-body tick)
  F_QSTATIC = 180,   // Gait::q_static_.head(7) (+1 pad)
  PLAN_STATE_DOUBLES = 188
};

struct PlanConst {
  double dt_mpc, dt_wbc, T_gait, T_mpc, h_ref, max_height, lock_time, lin_step;
  double shoulders[12];  // 3 x 4 column-major
  int N_gait, n_steps, k_mpc, pad;
};

// 0/1 matrix of up to 32 rows x 4 feet
struct GaitBits {
  unsigned long long lo, hi;
  QRW_HDB unsigned row(int r) const {
    if (r < 0 || r >= 32) return 0u;
    return (unsigned)((r < 16 ? lo >> (4 * r) : hi >> (4 * (r - 16))) & 15ull);
  }
  QRW_HDB void set(int r, unsigned v) {
    if (r < 16) lo = (lo & ~(15ull << (4 * r))) | ((unsigned long long)v << (4 * r));
    else hi = (hi & ~(15ull << (4 * (r - 16)))) | ((unsigned long long)v << (4 * (r - 16)));
  }
  QRW_HDB bool at(int r, int j) const { return (row(r) >> j) & 1u; }
  QRW_HDB void clear() { lo = hi = 0ull; }
  QRW_HDB void fill(int r0, int n, unsigned v, int N) {
    for (int r = r0; r < r0 + n && r < N; ++r) set(r, v);
  }
};

QRW_HDB int iround(double x) { return (int)lround(x); }

struct GaitState {
  GaitBits past, cur, des;
  int newPhase, is_static;
  double remaining;  // Gait::remainingTime_ (only meaningful right after phase_duration)

  // rows at or past N_gait read as zero, as the oracle's guard (the reference never gets there with N_gait = 20)
  QRW_HDB static bool rz(const GaitBits& m, int N, int r) { return r >= N || m.row(r) == 0u; }

  // Gait::create_pacing / bounding / trot / static / walk (Gait.cpp:39-109); foot j = bit j
  QRW_HDB void create(const PlanConst& C, int code) {
    const int N = C.N_gait;
    const int h = iround(0.5 * C.T_gait / C.dt_mpc);
    des.clear();
    if (code == 1) des.fill(0, h, 0x5u, N), des.fill(h, h, 0xAu, N);         // (1,0,1,0) then (0,1,0,1)
    else if (code == 2) des.fill(0, h, 0x3u, N), des.fill(h, h, 0xCu, N);    // (1,1,0,0) then (0,0,1,1)
    else if (code == 3) des.fill(0, h, 0x9u, N), des.fill(h, h, 0x6u, N);    // (1,0,0,1) then (0,1,1,0)
    else if (code == 4) des.fill(0, iround(C.T_gait / C.dt_mpc), 0xFu, N);   // static
    else if (code == 5) {                                                    // walk (private in the reference)
      const int q = iround(0.25 * C.T_gait / C.dt_mpc);
      des.fill(0, q, 0xEu, N), des.fill(q, q, 0xDu, N), des.fill(2 * q, q, 0xBu, N), des.fill(3 * q, q, 0x7u, N);
    }
  }
  // Gait::create_gait_f (Gait.cpp:111-140)
  QRW_HDB void create_gait_f(const PlanConst& C) {
    const int N = C.N_gait;
    int i = 0;
    for (int j = 0; j < C.n_steps; ++j) {
      cur.set(j, des.row(i));
      ++i;
      if (rz(des, N, i)) i = 0;
    }
    int index = 1;
    while (!rz(des, N, index)) ++index;
    for (int k = 0; k < i; ++k) rotate_up(des, index);
  }
  // rows 0..len-1 rotated up by one (row 0 goes to len-1): the swap loops of Gait.cpp:133-139, 240-258
  QRW_HDB static void rotate_up(GaitBits& m, int len) {
    const unsigned first = m.row(0);
    for (int r = 1; r < len; ++r) m.set(r - 1, m.row(r));
    m.set(len - 1, first);
  }
  // Gait::rollGait (Gait.cpp:223-260)
  QRW_HDB void roll(const PlanConst& C) {
    const int N = C.N_gait;
    const int top = C.n_steps < N ? C.n_steps : N - 1;  // the oracle's guard on row n_steps
    for (int m = top; m > 0; --m) past.set(m, past.row(m - 1));
    past.set(0, cur.row(0));
    newPhase = cur.row(0) != cur.row(1);
    int index = 1;
    while (!rz(cur, N, index)) ++index;
    rotate_up(cur, index);
    cur.set(index - 1, des.row(0));
    index = 1;
    while (!rz(des, N, index)) ++index;
    rotate_up(des, index);
  }
  // Gait::getPhaseDuration(i, j, value) (Gait.cpp:142-187); value = 1 stance, 0 swing
  QRW_HDB double phase_duration(const PlanConst& C, int i, int j, bool value) {
    const int N = C.N_gait;
    double t_phase = 1;
    int a = i;
    while (!rz(cur, N, i + 1) && cur.at(i + 1, j) == value) ++i, t_phase += 1;
    if (rz(cur, N, i + 1)) {
      int k = 0;
      while (!rz(des, N, k) && des.at(k, j) == value) ++k, t_phase += 1;
    }
    remaining = t_phase;
    while (a > 0 && cur.at(a - 1, j) == value) --a, t_phase += 1;
    if (a == 0)
      while (!rz(past, N, a) && past.at(a, j) == value) ++a, t_phase += 1;
    return t_phase * C.dt_mpc;
  }
};

// pinocchio::rpy::matrixToRpy(Eigen::Quaterniond(w, x, y, z).toRotationMatrix()), q = (.., .., .., x, y, z, w)
// (StatePlanner.cpp:23-24, FootstepPlanner.cpp:210-211); formula as restated in oracle/planner_oracle.cpp.
QRW_HD void quat_to_rpy(const double* q, double rpy[3]) {
  const double x = q[3], y = q[4], z = q[5], w = q[6];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y,
               tzz = tz * z;
  const double R00 = 1 - (tyy + tzz), R01 = txy - twz, R10 = txy + twz, R11 = 1 - (txx + tzz), R20 = txz - twy, R21 = tyz + twx,
               R22 = 1 - (txx + tyy);
  const double p = atan2(-R20, sqrt(R21 * R21 + R22 * R22));
  if (fabs(fabs(p) - 1.5707963267948966) < 0.001) {
    rpy[0] = 0.0, rpy[2] = -atan2(R01, R11);
  } else {
    rpy[2] = atan2(R10, R00), rpy[0] = atan2(R21, R22);
  }
  rpy[1] = p;
}

// sin and cos of one argument in one call (CUDA's sincos shares the argument reduction; same values as sin() / cos())
QRW_HD void sin_cos(double a, double& s, double& c) {
#if QRW_ON_GPU
  sincos(a, &s, &c);
#else
  s = sin(a), c = cos(a);
#endif
}

// One 5th-order swing polynomial (FootTrajectoryGenerator.cpp:58-71): the reference's closed forms
QRW_HD void swing_coeffs(double A[6], double a0, double v0, double x0, double xt, double t, double d) {
  const double t2 = t * t, t3 = t2 * t, t4 = t3 * t, t5 = t4 * t, d2 = d * d, d3 = d2 * d, d4 = d3 * d, d5 = d4 * d;
  const double cub = t3 - 3 * t2 * d + 3 * t * d2 - d3;
  const double den1 = 2 * ((t - d) * (t - d)) * cub;
  const double den2 = 2 * (t2 - 2 * t * d + d2) * cub;
  A[0] = (a0 * t2 - 2 * a0 * t * d - 6 * v0 * t + a0 * d2 + 6 * v0 * d + 12 * x0 - 12 * xt) / den1;
  A[1] = (30 * t * xt - 30 * t * x0 - 30 * d * x0 + 30 * d * xt - 2 * t3 * a0 - 3 * d3 * a0 + 14 * t2 * v0 - 16 * d2 * v0 + 2 * t * d * v0 +
          4 * t * d2 * a0 + t2 * d * a0) / den1;
  A[2] = (t4 * a0 + 3 * d4 * a0 - 8 * t3 * v0 + 12 * d3 * v0 + 20 * t2 * x0 - 20 * t2 * xt + 20 * d2 * x0 - 20 * d2 * xt + 80 * t * d * x0 -
          80 * t * d * xt + 4 * t3 * d * a0 + 28 * t * d2 * v0 - 32 * t2 * d * v0 - 8 * t2 * d2 * a0) / den1;
  A[3] = -(d5 * a0 + 4 * t * d4 * a0 + 3 * t4 * d * a0 + 36 * t * d3 * v0 - 24 * t3 * d * v0 + 60 * t * d2 * x0 + 60 * t2 * d * x0 -
           60 * t * d2 * xt - 60 * t2 * d * xt - 8 * t2 * d3 * a0 - 12 * t2 * d2 * v0) / den2;
  A[4] = -(2 * d5 * v0 - 2 * t * d5 * a0 - 10 * t * d4 * v0 + t2 * d4 * a0 + 4 * t3 * d3 * a0 - 3 * t4 * d2 * a0 - 16 * t2 * d3 * v0 +
           24 * t3 * d2 * v0 - 60 * t2 * d2 * x0 + 60 * t2 * d2 * xt) / den1;
  A[5] = (2 * xt * t5 - a0 * t4 * d3 - 10 * xt * t4 * d + 2 * a0 * t3 * d4 + 8 * v0 * t3 * d3 + 20 * xt * t3 * d2 - a0 * t2 * d5 -
          10 * v0 * t2 * d4 - 20 * x0 * t2 * d3 + 2 * v0 * t * d5 + 10 * x0 * t * d4 - 2 * x0 * d5) / den2;
}

// What one lane carries from `compute` to `store`
struct LaneCtx {
  GaitState g;
  int feet_mask;              // FootTrajectoryGenerator::feet as a bit mask
  int dirty;                  // stage bits whose per-foot state changed
  double cur_fs[3], target[3], o_target[3];
  double t0, tsw, ftg_target[3], Ax[6], Ay[6], pos[3], vel[3], acc[3];
  double pcmd[3], vcmd[3], acmd[3];
  double qstatic[7];
  int qstatic_set;
};

// Global-memory view of one planner object
struct PlanMem {
  int B;
  double* st;                 // [PLAN_STATE_DOUBLES][B] field-major
  unsigned long long* bits;   // [3][B][2] past / cur / des
  int* flags;                 // [B]: bit0 newPhase, bit1 is_static, bits 4..7 feet mask
  double* xref;               // [B][12*(n_steps+1)]
  double* fsteps;             // [B][N_gait*12]
  QRW_HD double* f(int field, int len, int b) const { return st + (size_t)B * field + (size_t)b * len; }
};

// Everything lane j of robot b computes in control tick k.  fs_col: where this lane's three footstep columns go
// (fs_col[i + N_gait*c], c = 0..2, i = 0..N_gait-1): shared-memory staging in the kernel, the output itself in emulation.
QRW_HD void lane_compute(const PlanConst& C, const PlanMem& M, int b, int j, int k, int stages, const double* q, const double* hv,
                         const double* vref, int code, double* fs_col, LaneCtx& X) {
  const int N = C.N_gait;
  const int B = M.B;
  GaitState& g = X.g;
  g.past.lo = M.bits[((size_t)0 * B + b) * 2], g.past.hi = M.bits[((size_t)0 * B + b) * 2 + 1];
  g.cur.lo = M.bits[((size_t)1 * B + b) * 2], g.cur.hi = M.bits[((size_t)1 * B + b) * 2 + 1];
  g.des.lo = M.bits[((size_t)2 * B + b) * 2], g.des.hi = M.bits[((size_t)2 * B + b) * 2 + 1];
  const int fl = M.flags[b];
  g.newPhase = fl & 1, g.is_static = (fl >> 1) & 1, g.remaining = 0.0;
  X.feet_mask = (fl >> 4) & 15;
  X.dirty = 0, X.qstatic_set = 0;

  // yaw (and roll / pitch for xref column 0) of the base, once for all the stages that need it
  double rpy[3] = {0.0, 0.0, 0.0}, cyaw = 1.0, syaw = 0.0;
  if (stages & (ST_FOOTSTEPS | ST_STATE | ST_FEETCMD)) {
    quat_to_rpy(q, rpy);
    sin_cos(rpy[2], syaw, cyaw);
  }

  // ---- Gait::updateGait (Gait.cpp:189-221)
  if (stages & ST_GAIT) {
    g.is_static = 0;
    if (code >= 1 && code <= 5) g.create(C, code);  // 1..4 as Gait::changeGait; 5 = create_walk (extension, see header)
    if (code == 4) {
      for (int i = 0; i < 7; ++i) X.qstatic[i] = q[i];
      X.qstatic_set = 1, g.is_static = 1;
    }
    if (k % C.k_mpc == 0) g.roll(C);
    X.dirty |= ST_GAIT;
  }

  // ---- FootstepPlanner::updateFootsteps (FootstepPlanner.cpp:52-75)
  if (stages & (ST_FOOTSTEPS | ST_FOOTTRAJ)) {
    const double* o = M.f(F_OTARGET, 12, b) + 3 * j;
    X.o_target[0] = o[0], X.o_target[1] = o[1], X.o_target[2] = o[2];
  }
  if (stages & ST_FOOTSTEPS) {
    const bool refresh = (k % C.k_mpc == 0) && k != 0;  // Controller.py:225
    const int kk = C.k_mpc - k % C.k_mpc;               // Controller.py:226
    const double* cf = M.f(F_CURFS, 12, b) + 3 * j;
    X.cur_fs[0] = cf[0], X.cur_fs[1] = cf[1], X.cur_fs[2] = cf[2];
    const double* tg = M.f(F_TARGET, 12, b) + 3 * j;
    X.target[0] = tg[0], X.target[1] = tg[1], X.target[2] = tg[2];
    const bool stance0 = g.cur.at(0, j);
    if (refresh && g.newPhase && stance0) {  // updateNewContact (:232-241): footsteps_[1] of the previous call
      const double* F = M.fsteps + (size_t)b * N * 12 + (size_t)N * 3 * j;
      X.cur_fs[0] = F[1], X.cur_fs[1] = F[1 + N], X.cur_fs[2] = F[1 + 2 * N];
    }
    if (stance0) {  // :62-73, the stance feet move backwards in the base frame
      const double rot = C.dt_wbc * vref[5];
      double c, s;
      sin_cos(rot, s, c);
      const double a = X.cur_fs[0] - C.dt_wbc * vref[0], bb = X.cur_fs[1] - C.dt_wbc * vref[1];
      X.cur_fs[0] = c * a + s * bb;
      X.cur_fs[1] = -s * a + c * bb;
    }
    // computeFootsteps (:77-154), this foot's columns only
    double prev[3] = {0.0, 0.0, 0.0};
    if (stance0) prev[0] = X.cur_fs[0], prev[1] = X.cur_fs[1], prev[2] = X.cur_fs[2];
    fs_col[0] = prev[0], fs_col[N] = prev[1], fs_col[2 * N] = prev[2];
    int first = (prev[0] != 0.0) ? 0 : -1;  // updateTargetFootsteps (:188-199): first row with a non-zero x
    double dt_cum = C.dt_wbc * kk;         // dt_cum(i-1) while row i is processed
    int i = 1;
    for (; !GaitState::rz(g.cur, N, i); ++i) {
      double v[3] = {0.0, 0.0, 0.0};
      const bool was = g.cur.at(i - 1, j), is = g.cur.at(i, j);
      if (was && is) {
        v[0] = prev[0], v[1] = prev[1], v[2] = prev[2];
      } else if (!was && is) {
        const double w = vref[5], yaw = w * dt_cum;
        double dx, dy;
        double c, s;  // yaws(i-1) = w * dt_cum(i-1) is also the argument of dx / dy (:96-111)
        sin_cos(yaw, s, c);
        if (w != 0) {
          dx = (hv[0] * s + hv[1] * (c - 1.0)) / w;
          dy = (hv[1] * s - hv[0] * (c - 1.0)) / w;
        } else {
          dx = hv[0] * dt_cum, dy = hv[1] * dt_cum;
        }
        // computeNextFootstep (:156-186)
        const double t_stance = g.phase_duration(C, i, j, true);
        double nx = t_stance * 0.5 * hv[0], ny = t_stance * 0.5 * hv[1];
        nx += 0.03 * (hv[0] - vref[0]), ny += 0.03 * (hv[1] - vref[1]);
        const double hg = 0.5 * sqrt(C.h_ref / 9.81);
        nx += hg * (hv[1] * vref[5] - hv[2] * vref[4]);
        ny += hg * (hv[2] * vref[3] - hv[0] * vref[5]);
        nx = dmax(dmin(nx, 0.155), -0.155), ny = dmax(dmin(ny, 0.155), -0.155);
        nx += C.shoulders[3 * j], ny += C.shoulders[3 * j + 1];
        v[0] = c * nx - s * ny + dx;
        v[1] = s * nx + c * ny + dy;
        v[2] = 0.0;
      }
      fs_col[i] = v[0], fs_col[i + N] = v[1], fs_col[i + 2 * N] = v[2];
      if (first < 0 && v[0] != 0.0) first = i, X.target[0] = v[0], X.target[1] = v[1], X.target[2] = 0.0;
      prev[0] = v[0], prev[1] = v[1], prev[2] = v[2];
      dt_cum = dt_cum + C.dt_mpc;
    }
    if (first == 0) X.target[0] = fs_col[0], X.target[1] = fs_col[N], X.target[2] = 0.0;
    for (; i < N; ++i) fs_col[i] = 0.0, fs_col[i + N] = 0.0, fs_col[i + 2 * N] = 0.0;
    // computeTargetFootstep (:201-230)
    const double c = cyaw, s = syaw;
    X.o_target[0] = c * X.target[0] - s * X.target[1] + q[0];
    X.o_target[1] = s * X.target[0] + c * X.target[1] + q[1];
    X.dirty |= ST_FOOTSTEPS;
  }

  // ---- FootTrajectoryGenerator::update (FootTrajectoryGenerator.cpp:110-151) with o_targetFootstep_ as its target
  if (stages & (ST_FOOTTRAJ | ST_FEETCMD)) {
    const double *p = M.f(F_FPOS, 12, b) + 3 * j, *v = M.f(F_FVEL, 12, b) + 3 * j, *a = M.f(F_FACC, 12, b) + 3 * j;
    for (int c = 0; c < 3; ++c) X.pos[c] = p[c], X.vel[c] = v[c], X.acc[c] = a[c];
  }
  if (stages & ST_FOOTTRAJ) {
    bool run = true;
    X.t0 = M.f(F_T0S, 4, b)[j], X.tsw = M.f(F_TSW, 4, b)[j];
    if ((k % C.k_mpc) == 0) {
      X.feet_mask = (int)(~g.cur.row(0) & 15u);
      if (X.feet_mask == 0) run = false;
      else if ((X.feet_mask >> j) & 1) {
        X.tsw = g.phase_duration(C, 0, j, false);
        const double value = X.tsw - (g.remaining * C.k_mpc - ((k + 1) % C.k_mpc)) * C.dt_wbc - C.dt_wbc;
        X.t0 = dmax(0.0, value);
      }
    } else {
      if (X.feet_mask == 0) run = false;
      else if ((X.feet_mask >> j) & 1) X.t0 = dmax(0.0, X.t0 + C.dt_wbc);
    }
    X.dirty |= ST_FOOTTRAJ;  // the feet mask is part of the state even when nothing else changes
    if (run && ((X.feet_mask >> j) & 1)) {
      // updateFootPosition (:41-108)
      const double t = X.t0, d = X.tsw;
      const double x0 = X.pos[0], y0 = X.pos[1];
      const double* ft = M.f(F_FTGTGT, 12, b) + 3 * j;
      X.ftg_target[0] = ft[0], X.ftg_target[1] = ft[1], X.ftg_target[2] = ft[2];
      const double *ax = M.f(F_AX, 24, b) + 6 * j, *ay = M.f(F_AY, 24, b) + 6 * j;
      for (int c = 0; c < 6; ++c) X.Ax[c] = ax[c], X.Ay[c] = ay[c];
      if (t < d - C.lock_time) {
        swing_coeffs(X.Ax, X.acc[0], X.vel[0], x0, X.o_target[0], t, d);
        swing_coeffs(X.Ay, X.acc[1], X.vel[1], y0, X.o_target[1], t, d);
        X.ftg_target[0] = X.o_target[0], X.ftg_target[1] = X.o_target[1];
      }
      const double hd = d / 2, gd = d - d / 2;
      const double den = (hd * hd * hd) * (gd * gd * gd);
      const double Az0 = -C.max_height / den, Az1 = (3 * d * C.max_height) / den, Az2 = -(3 * (d * d) * C.max_height) / den,
                   Az3 = ((d * d * d) * C.max_height) / den;
      const double ev = t + C.dt_wbc, e2 = ev * ev, e3 = e2 * ev, e4 = e3 * ev, e5 = e4 * ev, e6 = e5 * ev;
      if (t < 0.0 || t > d) {
        X.pos[0] = x0, X.pos[1] = y0, X.vel[0] = X.vel[1] = 0.0, X.acc[0] = X.acc[1] = 0.0;
      } else {
        const double* A = X.Ax;
        const double* Y = X.Ay;
        X.pos[0] = A[5] + A[4] * ev + A[3] * e2 + A[2] * e3 + A[1] * e4 + A[0] * e5;
        X.pos[1] = Y[5] + Y[4] * ev + Y[3] * e2 + Y[2] * e3 + Y[1] * e4 + Y[0] * e5;
        X.vel[0] = A[4] + 2 * A[3] * ev + 3 * A[2] * e2 + 4 * A[1] * e3 + 5 * A[0] * e4;
        X.vel[1] = Y[4] + 2 * Y[3] * ev + 3 * Y[2] * e2 + 4 * Y[1] * e3 + 5 * Y[0] * e4;
        X.acc[0] = 2 * A[3] + 3 * 2 * A[2] * ev + 4 * 3 * A[1] * e2 + 5 * 4 * A[0] * e3;
        X.acc[1] = 2 * Y[3] + 3 * 2 * Y[2] * ev + 4 * 3 * Y[1] * e2 + 5 * 4 * Y[0] * e3;
      }
      X.vel[2] = 3 * Az3 * e2 + 4 * Az2 * e3 + 5 * Az1 * e4 + 6 * Az0 * e5;
      X.acc[2] = 2 * 3 * Az3 * ev + 3 * 4 * Az2 * e2 + 4 * 5 * Az1 * e3 + 5 * 6 * Az0 * e4;
      X.pos[2] = Az3 * e3 + Az2 * e4 + Az1 * e5 + Az0 * e6;
      X.dirty |= 32;  // per-foot trajectory state changed
    }
  }

  // ---- StatePlanner::computeReferenceStates (StatePlanner.cpp:22-60), z_average = 0 (Controller.py:235-236);
  //      lane j writes columns j, j+4, ...
  if (stages & ST_STATE) {
    double* XR = M.xref + (size_t)b * 12 * (C.n_steps + 1);
    for (int col = j; col <= C.n_steps; col += 4) {
      double c12[12];
      for (int r = 0; r < 12; ++r) c12[r] = 0.0;
      if (col == 0) {
        c12[2] = q[2], c12[3] = rpy[0], c12[4] = rpy[1];
        for (int r = 0; r < 6; ++r) c12[6 + r] = hv[r];
      } else {
        const int i = col - 1;
        const double t = (i == C.n_steps - 1) ? C.T_mpc : C.dt_mpc + i * C.lin_step;  // Eigen LinSpaced(n, dt, T_mpc)
        const double w = vref[5];
        double sw, cw;  // referenceStates_(5, 1+i) = w * t is the argument of every sin / cos of the column
        sin_cos(w * t, sw, cw);
        if (w != 0) {
          c12[0] = (vref[0] * sw + vref[1] * (cw - 1.0)) / w;
          c12[1] = (vref[1] * sw - vref[0] * (cw - 1.0)) / w;
        } else {
          c12[0] = vref[0] * t, c12[1] = vref[1] * t;
        }
        c12[0] += 0.0, c12[1] += 0.0;  // + referenceStates_(0:2, 0), which are set to 0.0 (:27-28)
        c12[2] = C.h_ref + 0.0;
        c12[5] = w * t;
        c12[6] = vref[0] * cw - vref[1] * sw;
        c12[7] = vref[0] * sw + vref[1] * cw;
        c12[11] = w;
      }
      for (int r = 0; r < 12; ++r) XR[12 * col + r] = c12[r];
    }
  }

  // ---- feet commands in the base frame (Controller.py:286-296), oRh = Rz(yaw), oTh = (q_x, q_y, 0) (:419-423)
  if (stages & ST_FEETCMD) {
    const double *pc = M.f(F_PCMD, 12, b) + 3 * j, *vc = M.f(F_VCMD, 12, b) + 3 * j;
    const double c = cyaw, s = syaw;
    const double w[3] = {vref[3], vref[4], vref[5]};
    const double p0[3] = {pc[0], pc[1], pc[2]}, v0[3] = {vc[0], vc[1], vc[2]};
    const double wp[3] = {w[1] * p0[2] - w[2] * p0[1], w[2] * p0[0] - w[0] * p0[2], w[0] * p0[1] - w[1] * p0[0]};
    const double wwp[3] = {w[1] * wp[2] - w[2] * wp[1], w[2] * wp[0] - w[0] * wp[2], w[0] * wp[1] - w[1] * wp[0]};
    const double wv[3] = {w[1] * v0[2] - w[2] * v0[1], w[2] * v0[0] - w[0] * v0[2], w[0] * v0[1] - w[1] * v0[0]};
    // oRh' * u = (c ux + s uy, -s ux + c uy, uz)
    const double ra[3] = {c * X.acc[0] + s * X.acc[1], -s * X.acc[0] + c * X.acc[1], X.acc[2]};
    const double rv[3] = {c * X.vel[0] + s * X.vel[1], -s * X.vel[0] + c * X.vel[1], X.vel[2]};
    const double dp[3] = {X.pos[0] - 0.0 - q[0], X.pos[1] - 0.0 - q[1], X.pos[2] - C.h_ref - 0.0};
    for (int a = 0; a < 3; ++a) {
      X.acmd[a] = ra[a] - wwp[a] - 2 * wv[a];
      X.vcmd[a] = rv[a] - vref[a] - wp[a];
    }
    X.pcmd[0] = c * dp[0] + s * dp[1], X.pcmd[1] = -s * dp[0] + c * dp[1], X.pcmd[2] = dp[2];
    X.dirty |= ST_FEETCMD;
  }
}

// Write-back of what lane_compute changed (every lane's compute has finished by now)
QRW_HD void lane_store(const PlanConst& C, const PlanMem& M, int b, int j, const LaneCtx& X) {
  const int B = M.B;
  if (X.dirty & ST_GAIT) {
    M.f(F_CONTACTS, 4, b)[j] = X.g.cur.at(0, j) ? 1.0 : 0.0;
    if (j == 0) {
      M.bits[((size_t)0 * B + b) * 2] = X.g.past.lo, M.bits[((size_t)0 * B + b) * 2 + 1] = X.g.past.hi;
      M.bits[((size_t)1 * B + b) * 2] = X.g.cur.lo, M.bits[((size_t)1 * B + b) * 2 + 1] = X.g.cur.hi;
      M.bits[((size_t)2 * B + b) * 2] = X.g.des.lo, M.bits[((size_t)2 * B + b) * 2 + 1] = X.g.des.hi;
      if (X.qstatic_set)
        for (int i = 0; i < 7; ++i) M.f(F_QSTATIC, 8, b)[i] = X.qstatic[i];
    }
  }
  if (j == 0 && (X.dirty & (ST_GAIT | ST_FOOTTRAJ)))
    M.flags[b] = (X.g.newPhase & 1) | ((X.g.is_static & 1) << 1) | ((X.feet_mask & 15) << 4);
  if (X.dirty & ST_FOOTSTEPS) {
    double *cf = M.f(F_CURFS, 12, b) + 3 * j, *tg = M.f(F_TARGET, 12, b) + 3 * j, *o = M.f(F_OTARGET, 12, b) + 3 * j;
    for (int c = 0; c < 3; ++c) cf[c] = X.cur_fs[c], tg[c] = X.target[c];
    o[0] = X.o_target[0], o[1] = X.o_target[1];  // the z row of o_targetFootstep_ is never rewritten (:216-219)
  }
  if (X.dirty & 32) {
    M.f(F_T0S, 4, b)[j] = X.t0, M.f(F_TSW, 4, b)[j] = X.tsw;
    double *p = M.f(F_FPOS, 12, b) + 3 * j, *v = M.f(F_FVEL, 12, b) + 3 * j, *a = M.f(F_FACC, 12, b) + 3 * j,
           *ft = M.f(F_FTGTGT, 12, b) + 3 * j, *ax = M.f(F_AX, 24, b) + 6 * j, *ay = M.f(F_AY, 24, b) + 6 * j;
    for (int c = 0; c < 3; ++c) p[c] = X.pos[c], v[c] = X.vel[c], a[c] = X.acc[c], ft[c] = X.ftg_target[c];
    for (int c = 0; c < 6; ++c) ax[c] = X.Ax[c], ay[c] = X.Ay[c];
  } else if (X.dirty & ST_FOOTTRAJ) {
    M.f(F_T0S, 4, b)[j] = X.t0, M.f(F_TSW, 4, b)[j] = X.tsw;
  }
  if (X.dirty & ST_FEETCMD) {
    double *p = M.f(F_PCMD, 12, b) + 3 * j, *v = M.f(F_VCMD, 12, b) + 3 * j, *a = M.f(F_ACMD, 12, b) + 3 * j;
    for (int c = 0; c < 3; ++c) p[c] = X.pcmd[c], v[c] = X.vcmd[c], a[c] = X.acmd[c];
  }
}

}  // namespace plan
}  // namespace qrw
