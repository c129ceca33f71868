// planner_oracle.cpp -- CPU restatement of the reference's input producers.
//
// TEST INFRASTRUCTURE ONLY (see oracle/bindings.py): never linked, imported or executed by the product path.
//
// What is restated, function by function (all citations into /root/reference):
//   Gait                     src/Gait.cpp:3-269, include/qrw/Gait.hpp
//   StatePlanner             src/StatePlanner.cpp:3-61
//   FootstepPlanner          src/FootstepPlanner.cpp:3-249
//   FootTrajectoryGenerator  src/FootTrajectoryGenerator.cpp:5-151
//   and the order in which scripts/Controller.py:222-240 calls them in one control tick.
//
// Matrices are plain double arrays with the reference's shapes; Eigen's isZero() (precision 1e-12) and isApprox()
// are restated for the 0/1 rows they are applied to.  The only third-party arithmetic on this path is
// pinocchio::rpy::matrixToRpy (StatePlanner.cpp:24, FootstepPlanner.cpp:211); Pinocchio is an unpinned dependency
// that is absent here, so its published formula (pinocchio/math/rpy.hxx, v2.5+: pitch = atan2(-R20, hypot(R21, R22)),
// yaw = atan2(R10, R00), roll = atan2(R21, R22), with the gimbal-lock branch within 1e-3 of +-pi/2) is restated.
// PARITY UNPINNED: the reference has no test or golden vector for these classes, and none of them can be built here
// (Eigen and Pinocchio headers are missing); the restatement is pinned by the properties in tests/test_planner_oracle.py.
//
// Places where the reference reads out of bounds are guarded and documented at the guard (they cannot be reached with
// the reference's own configuration N_gait = 20, n_steps = 16).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace {

constexpr int MAXG = 32;  // N_gait <= 32, like the MPC side of the ABI

struct Gait {
  int N = 0, n_steps = 0;
  double dt = 0, T_gait = 0, T_mpc = 0, remainingTime = 0;
  bool newPhase = false, is_static = true;
  double q_static[19] = {0};
  double past[MAXG][4], cur[MAXG][4], des[MAXG][4];

  static bool rowZero(const double (*m)[4], int N, int r) {  // Eigen isZero(1e-12); rows past the end read as zero
    if (r < 0 || r >= N) return true;
    for (int j = 0; j < 4; ++j)
      if (std::fabs(m[r][j]) > 1e-12) return false;
    return true;
  }
  static void zero(double (*m)[4]) { std::memset(m, 0, sizeof(double) * MAXG * 4); }
  static void swapRows(double (*m)[4], int a, int b) {
    for (int j = 0; j < 4; ++j) std::swap(m[a][j], m[b][j]);
  }
  void fill(int r0, int n, double a, double b, double c, double d) {
    for (int r = r0; r < r0 + n && r < N; ++r) des[r][0] = a, des[r][1] = b, des[r][2] = c, des[r][3] = d;
  }
  // Gait.cpp:21-36 (returns false where the reference throws std::invalid_argument)
  bool initialize(double dt_in, double T_gait_in, double T_mpc_in, int N_gait) {
    dt = dt_in, T_gait = T_gait_in, T_mpc = T_mpc_in, N = N_gait;
    n_steps = (int)std::lround(T_mpc_in / dt_in);
    zero(past), zero(cur), zero(des);
    if (n_steps > N_gait || (int)std::lround(T_gait_in / dt_in) > N_gait) return false;
    create_trot();
    create_gait_f();
    return true;
  }
  void create_walk() {  // Gait.cpp:39-55
    int n = (int)std::lround(0.25 * T_gait / dt);
    zero(des);
    fill(0, n, 0, 1, 1, 1), fill(n, n, 1, 0, 1, 1), fill(2 * n, n, 1, 1, 0, 1), fill(3 * n, n, 1, 1, 1, 0);
  }
  void create_trot() {  // Gait.cpp:57-69
    int n = (int)std::lround(0.5 * T_gait / dt);
    zero(des);
    fill(0, n, 1, 0, 0, 1), fill(n, n, 0, 1, 1, 0);
  }
  void create_pacing() {  // Gait.cpp:71-83
    int n = (int)std::lround(0.5 * T_gait / dt);
    zero(des);
    fill(0, n, 1, 0, 1, 0), fill(n, n, 0, 1, 0, 1);
  }
  void create_bounding() {  // Gait.cpp:85-97
    int n = (int)std::lround(0.5 * T_gait / dt);
    zero(des);
    fill(0, n, 1, 1, 0, 0), fill(n, n, 0, 0, 1, 1);
  }
  void create_static() {  // Gait.cpp:99-109
    int n = (int)std::lround(T_gait / dt);
    zero(des);
    fill(0, n, 1, 1, 1, 1);
  }
  void create_gait_f() {  // Gait.cpp:111-140
    int i = 0;
    for (int j = 0; j < n_steps; ++j) {
      for (int c = 0; c < 4; ++c) cur[j][c] = des[i][c];
      ++i;
      if (rowZero(des, N, i)) i = 0;
    }
    int index = 1;
    while (!rowZero(des, N, index)) ++index;
    for (int k = 0; k < i; ++k)
      for (int m = 0; m < index - 1; ++m) swapRows(des, m, m + 1);
  }
  // Gait.cpp:142-187
  double getPhaseDuration(int i, int j, double value) {
    double t_phase = 1;
    int a = i;
    while (!rowZero(cur, N, i + 1) && cur[i + 1][j] == value) ++i, ++t_phase;
    if (rowZero(cur, N, i + 1)) {
      int k = 0;
      while (!rowZero(des, N, k) && des[k][j] == value) ++k, ++t_phase;
    }
    remainingTime = t_phase;
    while (a > 0 && cur[a - 1][j] == value) --a, ++t_phase;
    if (a == 0)
      while (!rowZero(past, N, a) && past[a][j] == value) ++a, ++t_phase;
    return t_phase * dt;
  }
  // Gait.cpp:199-221
  bool changeGait(int code, const double* q) {
    is_static = false;
    if (code == 1) create_pacing();
    else if (code == 2) create_bounding();
    else if (code == 3) create_trot();
    else if (code == 4) {
      create_static();
      for (int i = 0; i < 7; ++i) q_static[i] = q[i];
      is_static = true;
    } else if (code == 5) {
      create_walk();  // EXTENSION (not in Gait::changeGait): the reference's private, never-called create_walk
    }
    return is_static;
  }
  // Gait.cpp:223-260
  void rollGait() {
    for (int m = n_steps; m > 0; --m)
      if (m < N) swapRows(past, m, m - 1);  // guard: row n_steps exists only when N_gait > n_steps
    for (int c = 0; c < 4; ++c) past[0][c] = cur[0][c];
    bool same = true;  // isApprox of two 0/1 rows == equality
    for (int c = 0; c < 4; ++c) same = same && cur[0][c] == cur[1][c];
    newPhase = !same;
    int index = 1;
    while (!rowZero(cur, N, index)) swapRows(cur, index - 1, index), ++index;
    for (int c = 0; c < 4; ++c) cur[index - 1][c] = des[0][c];
    index = 1;
    while (!rowZero(des, N, index)) swapRows(des, index - 1, index), ++index;
  }
  void updateGait(int k, int k_mpc, const double* q, int code) {  // Gait.cpp:189-197
    changeGait(code, q);
    if (k % k_mpc == 0) rollGait();
  }
};

// pinocchio::rpy::matrixToRpy of the rotation of quaternion (w, x, y, z) = (q[6], q[3], q[4], q[5])
void quatToRpy(const double* q, double rpy[3]) {
  const double x = q[3], y = q[4], z = q[5], w = q[6];
  // Eigen::Quaterniond::toRotationMatrix (no normalisation)
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y,
               tyz = tz * y, tzz = tz * z;
  const double R00 = 1 - (tyy + tzz), R01 = txy - twz, R10 = txy + twz, R11 = 1 - (txx + tzz), R20 = txz - twy,
               R21 = tyz + twx, R22 = 1 - (txx + tyy);
  const double m = std::sqrt(R21 * R21 + R22 * R22);
  const double p = std::atan2(-R20, m);
  double r, yw;
  if (std::fabs(std::fabs(p) - M_PI / 2) < 0.001) {
    r = 0.0;
    yw = -std::atan2(R01, R11);
  } else {
    yw = std::atan2(R10, R00);
    r = std::atan2(R21, R22);
  }
  rpy[0] = r, rpy[1] = p, rpy[2] = yw;
}

struct StatePlanner {  // src/StatePlanner.cpp
  double dt = 0, h_ref = 0;
  int n_steps = 0;
  std::vector<double> xref, dtv;  // 12 x (1+n) column-major; Eigen LinSpaced(n, dt, T_mpc)
  void initialize(double dt_in, double T_mpc_in, double h_ref_in) {  // :13-20
    dt = dt_in, h_ref = h_ref_in;
    n_steps = (int)std::lround(T_mpc_in / dt_in);
    xref.assign(12 * (1 + n_steps), 0.0);
    dtv.resize(n_steps);
    // Eigen::VectorXd::LinSpaced(size, low, high): low + i*step, the last element is `high` itself
    const double step = n_steps > 1 ? (T_mpc_in - dt_in) / (n_steps - 1) : 0.0;
    for (int i = 0; i < n_steps; ++i) dtv[i] = (i == n_steps - 1) ? T_mpc_in : dt_in + i * step;
  }
  void compute(const double* q, const double* v, const double* vref, double z_average) {  // :22-60
    double rpy[3];
    quatToRpy(q, rpy);
    double* X = xref.data();
    X[0] = 0.0, X[1] = 0.0, X[2] = q[2], X[3] = rpy[0], X[4] = rpy[1], X[5] = 0.0;
    for (int i = 0; i < 3; ++i) X[6 + i] = v[i], X[9 + i] = v[3 + i];
    for (int i = 0; i < n_steps; ++i) {
      double* c = X + 12 * (1 + i);
      const double t = dtv[i];
      if (vref[5] != 0) {
        c[0] = (vref[0] * std::sin(vref[5] * t) + vref[1] * (std::cos(vref[5] * t) - 1.0)) / vref[5];
        c[1] = (vref[1] * std::sin(vref[5] * t) - vref[0] * (std::cos(vref[5] * t) - 1.0)) / vref[5];
      } else {
        c[0] = vref[0] * t;
        c[1] = vref[1] * t;
      }
      c[0] += X[0];
      c[1] += X[1];
      c[2] = h_ref + z_average;
      c[5] = vref[5] * t;
      c[6] = vref[0] * std::cos(c[5]) - vref[1] * std::sin(c[5]);
      c[7] = vref[0] * std::sin(c[5]) + vref[1] * std::cos(c[5]);
      c[11] = vref[5];
    }
  }
};

struct FootstepPlanner {  // src/FootstepPlanner.cpp
  Gait* gait = nullptr;
  double dt = 0, dt_wbc = 0, T_mpc = 0, h_ref = 0, k_feedback = 0.03, g = 9.81, L = 0.155;
  int n_steps = 0, N = 0;
  double shoulders[4][3], cur[4][3], next[4][3], target[4][3], o_target[4][3];  // [foot][xyz] = columns of Matrix34
  std::vector<double> fs;  // footsteps_: [row][foot][xyz]
  double dt_cum[MAXG], yaws[MAXG], dx[MAXG], dy[MAXG];
  double* F(int i, int j) { return fs.data() + (i * 4 + j) * 3; }

  void initialize(double dt_in, double dt_wbc_in, double T_mpc_in, double h_ref_in, const double* sh34, Gait& g_, int N_gait) {  // :23-50
    dt = dt_in, dt_wbc = dt_wbc_in, T_mpc = T_mpc_in, h_ref = h_ref_in, N = N_gait;
    n_steps = (int)std::lround(T_mpc_in / dt_in);
    for (int j = 0; j < 4; ++j)
      for (int c = 0; c < 3; ++c) shoulders[j][c] = cur[j][c] = target[j][c] = o_target[j][c] = sh34[c + 3 * j], next[j][c] = 0;
    gait = &g_;
    fs.assign((size_t)N * 12, 0.0);
    std::fill(dt_cum, dt_cum + MAXG, 0.0), std::fill(yaws, yaws + MAXG, 0.0), std::fill(dx, dx + MAXG, 0.0), std::fill(dy, dy + MAXG, 0.0);
  }
  void updateNewContact() {  // :232-241
    for (int i = 0; i < 4; ++i)
      if (gait->cur[0][i] == 1.0)
        for (int c = 0; c < 3; ++c) cur[i][c] = F(1, i)[c];
  }
  void computeNextFootstep(int i, int j, const double* b_v, const double* b_vref) {  // :156-186
    std::memset(next, 0, sizeof(next));
    const double t_stance = gait->getPhaseDuration(i, j, 1.0);
    for (int c = 0; c < 3; ++c) next[j][c] = t_stance * 0.5 * b_v[c];
    for (int c = 0; c < 3; ++c) next[j][c] += k_feedback * (b_v[c] - b_vref[c]);
    const double cross[3] = {b_v[1] * b_vref[5] - b_v[2] * b_vref[4], b_v[2] * b_vref[3] - b_v[0] * b_vref[5], 0.0};
    for (int c = 0; c < 3; ++c) next[j][c] += 0.5 * std::sqrt(h_ref / g) * cross[c];
    next[j][0] = std::min(next[j][0], L), next[j][0] = std::max(next[j][0], -L);
    next[j][1] = std::min(next[j][1], L), next[j][1] = std::max(next[j][1], -L);
    for (int c = 0; c < 3; ++c) next[j][c] += shoulders[j][c];
    for (int f = 0; f < 4; ++f) next[f][2] = 0.0;
  }
  void computeFootsteps(int k, const double* b_v, const double* b_vref) {  // :77-154
    std::fill(fs.begin(), fs.end(), 0.0);
    const double(*G)[4] = gait->cur;
    for (int j = 0; j < 4; ++j)
      if (G[0][j] == 1.0)
        for (int c = 0; c < 3; ++c) F(0, j)[c] = cur[j][c];
    dt_cum[0] = dt_wbc * k;
    yaws[0] = b_vref[5] * dt_cum[0];
    for (int j = 1; j < N; ++j) {
      dt_cum[j] = Gait::rowZero(G, N, j) ? dt_cum[j - 1] : dt_cum[j - 1] + dt;
      yaws[j] = b_vref[5] * dt_cum[j];
    }
    if (b_vref[5] != 0) {
      for (int j = 0; j < N; ++j) {
        dx[j] = (b_v[0] * std::sin(b_vref[5] * dt_cum[j]) + b_v[1] * (std::cos(b_vref[5] * dt_cum[j]) - 1.0)) / b_vref[5];
        dy[j] = (b_v[1] * std::sin(b_vref[5] * dt_cum[j]) - b_v[0] * (std::cos(b_vref[5] * dt_cum[j]) - 1.0)) / b_vref[5];
      }
    } else {
      for (int j = 0; j < N; ++j) dx[j] = b_v[0] * dt_cum[j], dy[j] = b_v[1] * dt_cum[j];
    }
    int i = 1;
    while (!Gait::rowZero(G, N, i)) {
      for (int j = 0; j < 4; ++j)
        if (G[i - 1][j] * G[i][j] > 0)
          for (int c = 0; c < 3; ++c) F(i, j)[c] = F(i - 1, j)[c];
      for (int j = 0; j < 4; ++j)
        if ((1 - G[i - 1][j]) * G[i][j] > 0) {
          const double qd[3] = {dx[i - 1], dy[i - 1], 0.0};
          computeNextFootstep(i, j, b_v, b_vref);
          const double c = std::cos(yaws[i - 1]), s = std::sin(yaws[i - 1]);
          F(i, j)[0] = c * next[j][0] - s * next[j][1] + qd[0];
          F(i, j)[1] = s * next[j][0] + c * next[j][1] + qd[1];
          F(i, j)[2] = next[j][2] + qd[2];
        }
      ++i;
    }
  }
  void updateTargetFootsteps() {  // :188-199
    for (int i = 0; i < 4; ++i) {
      int index = 0;
      while (index < N && F(index, i)[0] == 0.0) ++index;
      if (index == N) continue;  // guard: the reference reads past the end when a foot never lands in the horizon
      target[i][0] = F(index, i)[0], target[i][1] = F(index, i)[1], target[i][2] = 0.0;
    }
  }
  const double* computeTargetFootstep(int k, const double* q, const double* b_v, const double* b_vref) {  // :201-230
    computeFootsteps(k, b_v, b_vref);
    updateTargetFootsteps();
    double rpy[3];
    quatToRpy(q, rpy);
    const double c = std::cos(rpy[2]), s = std::sin(rpy[2]);
    for (int i = 0; i < 4; ++i) {
      o_target[i][0] = c * target[i][0] - s * target[i][1] + q[0];
      o_target[i][1] = s * target[i][0] + c * target[i][1] + q[1];
    }
    return &o_target[0][0];
  }
  const double* updateFootsteps(bool refresh, int k, const double* q, const double* b_v, const double* b_vref) {  // :52-75
    if (refresh && gait->newPhase) updateNewContact();
    // :62-73.  The reference multiplies the 3x3 `Rz` by a 2x1 block (:71); compiled with -DNDEBUG the product's first
    // two rows are [c s; -s c] * (p - dpos) + 0 * (whatever lies behind the 2-vector), i.e. the planar rotation
    // restated here (Rz(0,2) = Rz(1,2) = 0 always).
    const double rot = dt_wbc * b_vref[5];
    const double c = std::cos(rot), s = std::sin(rot);
    const double dpos[2] = {dt_wbc * b_vref[0], dt_wbc * b_vref[1]};
    for (int j = 0; j < 4; ++j)
      if (gait->cur[0][j] == 1.0) {
        const double a = cur[j][0] - dpos[0], b = cur[j][1] - dpos[1];
        cur[j][0] = c * a + s * b;
        cur[j][1] = -s * a + c * b;
      }
    return computeTargetFootstep(k, q, b_v, b_vref);
  }
};

struct FootTrajectoryGenerator {  // src/FootTrajectoryGenerator.cpp
  Gait* gait = nullptr;
  double dt_tsid = 0, maxHeight = 0, lockTime = 0;
  int k_mpc = 0;
  std::vector<int> feet;
  double t0s[4] = {0}, t_swing[4] = {0};
  double target[4][3] = {{0}}, Ax[4][6] = {{0}}, Ay[4][6] = {{0}}, pos[4][3] = {{0}}, vel[4][3] = {{0}}, acc[4][3] = {{0}};

  void initialize(double maxHeightIn, double lockTimeIn, const double* target34, const double* pos34, double dt_tsid_in, int k_mpc_in, Gait& g_) {  // :23-38
    dt_tsid = dt_tsid_in, k_mpc = k_mpc_in, maxHeight = maxHeightIn, lockTime = lockTimeIn, gait = &g_;
    for (int j = 0; j < 4; ++j)
      for (int c = 0; c < 3; ++c) target[j][c] = target34[c + 3 * j], pos[j][c] = pos34[c + 3 * j];
  }
  // :41-108.  The twelve coefficient formulas are the reference's closed forms with std::pow(x, n) written as
  // products (pow with a small integer exponent is exact repeated multiplication up to the last ulp).
  static void coeffs(double A[6], double a0, double v0, double x0, double xt, double t, double d) {
    const double t2 = t * t, t3 = t2 * t, t4 = t3 * t, t5 = t4 * t, d2 = d * d, d3 = d2 * d, d4 = d3 * d, d5 = d4 * d;
    const double cub = t3 - 3 * t2 * d + 3 * t * d2 - d3;
    const double den1 = 2 * ((t - d) * (t - d)) * cub;
    const double den2 = 2 * (t2 - 2 * t * d + d2) * cub;
    A[0] = (a0 * t2 - 2 * a0 * t * d - 6 * v0 * t + a0 * d2 + 6 * v0 * d + 12 * x0 - 12 * xt) / den1;
    A[1] = (30 * t * xt - 30 * t * x0 - 30 * d * x0 + 30 * d * xt - 2 * t3 * a0 - 3 * d3 * a0 + 14 * t2 * v0 - 16 * d2 * v0 +
            2 * t * d * v0 + 4 * t * d2 * a0 + t2 * d * a0) / den1;
    A[2] = (t4 * a0 + 3 * d4 * a0 - 8 * t3 * v0 + 12 * d3 * v0 + 20 * t2 * x0 - 20 * t2 * xt + 20 * d2 * x0 - 20 * d2 * xt +
            80 * t * d * x0 - 80 * t * d * xt + 4 * t3 * d * a0 + 28 * t * d2 * v0 - 32 * t2 * d * v0 - 8 * t2 * d2 * a0) / den1;
    A[3] = -(d5 * a0 + 4 * t * d4 * a0 + 3 * t4 * d * a0 + 36 * t * d3 * v0 - 24 * t3 * d * v0 + 60 * t * d2 * x0 + 60 * t2 * d * x0 -
             60 * t * d2 * xt - 60 * t2 * d * xt - 8 * t2 * d3 * a0 - 12 * t2 * d2 * v0) / den2;
    A[4] = -(2 * d5 * v0 - 2 * t * d5 * a0 - 10 * t * d4 * v0 + t2 * d4 * a0 + 4 * t3 * d3 * a0 - 3 * t4 * d2 * a0 - 16 * t2 * d3 * v0 +
             24 * t3 * d2 * v0 - 60 * t2 * d2 * x0 + 60 * t2 * d2 * xt) / den1;
    A[5] = (2 * xt * t5 - a0 * t4 * d3 - 10 * xt * t4 * d + 2 * a0 * t3 * d4 + 8 * v0 * t3 * d3 + 20 * xt * t3 * d2 - a0 * t2 * d5 -
            10 * v0 * t2 * d4 - 20 * x0 * t2 * d3 + 2 * v0 * t * d5 + 10 * x0 * t * d4 - 2 * x0 * d5) / den2;
  }
  void updateFootPosition(int j, const double* tgt) {
    const double ddx0 = acc[j][0], ddy0 = acc[j][1], dx0 = vel[j][0], dy0 = vel[j][1], x0 = pos[j][0], y0 = pos[j][1];
    const double t = t0s[j], d = t_swing[j], dt = dt_tsid;
    if (t < d - lockTime) {
      coeffs(Ax[j], ddx0, dx0, x0, tgt[0], t, d);
      coeffs(Ay[j], ddy0, dy0, y0, tgt[1], t, d);
      target[j][0] = tgt[0], target[j][1] = tgt[1];
    }
    const double h3 = (d / 2) * (d / 2) * (d / 2), g3 = (d - d / 2) * (d - d / 2) * (d - d / 2);
    const double Az0 = -maxHeight / (h3 * g3), Az1 = (3 * d * maxHeight) / (h3 * g3), Az2 = -(3 * (d * d) * maxHeight) / (h3 * g3),
                 Az3 = ((d * d * d) * maxHeight) / (h3 * g3);
    const double ev = t + dt, e2 = ev * ev, e3 = e2 * ev, e4 = e3 * ev, e5 = e4 * ev, e6 = e5 * ev;
    if (t < 0.0 || t > d) {
      pos[j][0] = x0, pos[j][1] = y0, vel[j][0] = vel[j][1] = 0.0, acc[j][0] = acc[j][1] = 0.0;
    } else {
      const double* X = Ax[j];
      const double* Y = Ay[j];
      pos[j][0] = X[5] + X[4] * ev + X[3] * e2 + X[2] * e3 + X[1] * e4 + X[0] * e5;
      pos[j][1] = Y[5] + Y[4] * ev + Y[3] * e2 + Y[2] * e3 + Y[1] * e4 + Y[0] * e5;
      vel[j][0] = X[4] + 2 * X[3] * ev + 3 * X[2] * e2 + 4 * X[1] * e3 + 5 * X[0] * e4;
      vel[j][1] = Y[4] + 2 * Y[3] * ev + 3 * Y[2] * e2 + 4 * Y[1] * e3 + 5 * Y[0] * e4;
      acc[j][0] = 2 * X[3] + 3 * 2 * X[2] * ev + 4 * 3 * X[1] * e2 + 5 * 4 * X[0] * e3;
      acc[j][1] = 2 * Y[3] + 3 * 2 * Y[2] * ev + 4 * 3 * Y[1] * e2 + 5 * 4 * Y[0] * e3;
    }
    vel[j][2] = 3 * Az3 * e2 + 4 * Az2 * e3 + 5 * Az1 * e4 + 6 * Az0 * e5;
    acc[j][2] = 2 * 3 * Az3 * ev + 3 * 4 * Az2 * e2 + 4 * 5 * Az1 * e3 + 5 * 6 * Az0 * e4;
    pos[j][2] = Az3 * e3 + Az2 * e4 + Az1 * e5 + Az0 * e6;
  }
  void update(int k, const double* tgt34) {  // :110-151
    if ((k % k_mpc) == 0) {
      feet.clear();
      for (int i = 0; i < 4; ++i)
        if (gait->cur[0][i] == 0) feet.push_back(i);
      if (feet.empty()) return;
      for (int i : feet) {
        t_swing[i] = gait->getPhaseDuration(0, i, 0.0);
        const double value = t_swing[i] - (gait->remainingTime * k_mpc - ((k + 1) % k_mpc)) * dt_tsid - dt_tsid;
        t0s[i] = std::max(0.0, value);
      }
    } else {
      if (feet.empty()) return;
      for (int i : feet) t0s[i] = std::max(0.0, t0s[i] + dt_tsid);
    }
    for (int i : feet) updateFootPosition(i, tgt34 + 3 * i);
  }
};

// One robot's planners wired as scripts/Controller.py:124-139 does.
struct Robot {
  Gait gait;
  StatePlanner sp;
  FootstepPlanner fsp;
  FootTrajectoryGenerator ftg;
};

struct Planner {
  int B, N_gait, k_mpc, n_steps;
  std::vector<Robot> r;
};

}  // namespace

// stage bits of orc_planner_run (the same as qrw_planner_run_*): one bit per reference call of Controller.py:222-236
enum { ST_GAIT = 1, ST_FOOTSTEPS = 2, ST_FOOTTRAJ = 4, ST_STATE = 8 };

extern "C" {

// Gait::initialize + StatePlanner::initialize + FootstepPlanner::initialize + FootTrajectoryGenerator::initialize with
// the arguments of Controller.py:124-139.  shoulders / ftg_target / ftg_pos: 3 x 4 column-major, shared by the batch.
void* orc_planner_create(int batch, double dt_mpc, double dt_wbc, double T_gait, double T_mpc, int N_gait, int k_mpc, double h_ref,
                         const double* shoulders, const double* ftg_target, const double* ftg_pos, double max_height,
                         double lock_time) {
  if (N_gait > MAXG || batch <= 0) return nullptr;
  auto* P = new Planner();
  P->B = batch, P->N_gait = N_gait, P->k_mpc = k_mpc;
  P->r.resize(batch);
  for (auto& R : P->r) {
    if (!R.gait.initialize(dt_mpc, T_gait, T_mpc, N_gait)) {
      delete P;
      return nullptr;
    }
    R.sp.initialize(dt_mpc, T_mpc, h_ref);
    R.fsp.initialize(dt_mpc, dt_wbc, T_mpc, h_ref, shoulders, R.gait, N_gait);
    R.ftg.initialize(max_height, lock_time, ftg_target, ftg_pos, dt_wbc, k_mpc, R.gait);
  }
  P->n_steps = P->r[0].sp.n_steps;
  return P;
}
void orc_planner_destroy(void* h) { delete static_cast<Planner*>(h); }

// One control tick k of every robot.  q7 [B][7] (x y z, quaternion x y z w), v6 = h_v [B][6], vref6 [B][6], code [B]
// (joystick code of Gait::changeGait; NULL = 0 for all).
void orc_planner_run(void* h, int k, int stages, const double* q7, const double* v6, const double* vref6, const int* code) {
  auto* P = static_cast<Planner*>(h);
  for (int b = 0; b < P->B; ++b) {
    Robot& R = P->r[b];
    const double *q = q7 + 7 * b, *v = v6 + 6 * b, *vr = vref6 + 6 * b;
    if (stages & ST_GAIT) R.gait.updateGait(k, P->k_mpc, q, code ? code[b] : 0);
    if (stages & ST_FOOTSTEPS)
      R.fsp.updateFootsteps(k % P->k_mpc == 0 && k != 0, P->k_mpc - k % P->k_mpc, q, v, vr);  // Controller.py:225-229
    if (stages & ST_FOOTTRAJ) R.ftg.update(k, &R.fsp.o_target[0][0]);
    if (stages & ST_STATE) R.sp.compute(q, v, vr, 0.0);
  }
}

// Outputs in the C-ABI layout (instance-major, column-major inside): xref [B][12*(n+1)], fsteps [B][N_gait*12]
// (FootstepPlanner::getFootsteps = vectorToMatrix, N_gait x 12), gait [B][N_gait*4] (getCurrentGait), o_target, fpos,
// fvel, facc [B][12] (3 x 4).  Any pointer may be NULL.
void orc_planner_outputs(void* h, double* xref, double* fsteps, double* gait, double* o_target, double* fpos, double* fvel,
                         double* facc) {
  auto* P = static_cast<Planner*>(h);
  const int N = P->N_gait, n1 = P->n_steps + 1;
  for (int b = 0; b < P->B; ++b) {
    Robot& R = P->r[b];
    if (xref) std::copy(R.sp.xref.begin(), R.sp.xref.end(), xref + (size_t)b * 12 * n1);
    if (fsteps)
      for (int i = 0; i < N; ++i)
        for (int j = 0; j < 4; ++j)
          for (int c = 0; c < 3; ++c) fsteps[(size_t)b * N * 12 + i + N * (3 * j + c)] = R.fsp.F(i, j)[c];
    if (gait)
      for (int i = 0; i < N; ++i)
        for (int j = 0; j < 4; ++j) gait[(size_t)b * N * 4 + i + N * j] = R.gait.cur[i][j];
    for (int j = 0; j < 4; ++j)
      for (int c = 0; c < 3; ++c) {
        const size_t o = (size_t)b * 12 + 3 * j + c;
        if (o_target) o_target[o] = R.fsp.o_target[j][c];
        if (fpos) fpos[o] = R.ftg.pos[j][c];
        if (fvel) fvel[o] = R.ftg.vel[j][c];
        if (facc) facc[o] = R.ftg.acc[j][c];
      }
  }
}

// Internal state for the tests: past / desired gait [B][N_gait*4], current footsteps [B][12], t0s / t_swing [B][4],
// flags [B][2] = (newPhase, is_static).  Any pointer may be NULL.
void orc_planner_state(void* h, double* past, double* desired, double* cur_fs, double* t0s, double* t_swing, int* flags) {
  auto* P = static_cast<Planner*>(h);
  const int N = P->N_gait;
  for (int b = 0; b < P->B; ++b) {
    Robot& R = P->r[b];
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < 4; ++j) {
        if (past) past[(size_t)b * N * 4 + i + N * j] = R.gait.past[i][j];
        if (desired) desired[(size_t)b * N * 4 + i + N * j] = R.gait.des[i][j];
      }
    for (int j = 0; j < 4; ++j) {
      for (int c = 0; c < 3; ++c)
        if (cur_fs) cur_fs[(size_t)b * 12 + 3 * j + c] = R.fsp.cur[j][c];
      if (t0s) t0s[4 * b + j] = R.ftg.t0s[j];
      if (t_swing) t_swing[4 * b + j] = R.ftg.t_swing[j];
    }
    if (flags) flags[2 * b] = R.gait.newPhase, flags[2 * b + 1] = R.gait.is_static;
  }
}

// Gait::getPhaseDuration(i, j, value) of robot b (also returns remainingTime_ through *remaining)
double orc_planner_phase_duration(void* h, int b, int i, int j, double value, double* remaining) {
  auto* P = static_cast<Planner*>(h);
  const double t = P->r[b].gait.getPhaseDuration(i, j, value);
  if (remaining) *remaining = P->r[b].gait.remainingTime;
  return t;
}
// desired gait of one robot set to "walk" (Gait::create_walk, Gait.cpp:39-55; private and never called in the
// reference, exposed here because BASELINE config 3 names a walk schedule)
void orc_planner_create_walk(void* h, int b) { static_cast<Planner*>(h)->r[b].gait.create_walk(); }

}  // extern "C"
