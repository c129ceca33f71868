// oracle/qrw_oracle.cpp
//
// TEST INFRASTRUCTURE ONLY -- CPU oracle for the qrw per-tick hot path.  Not part of the
// product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, load or call this library.
//
// A dependency-free C++17 restatement of
//   * MPC     -- /root/reference/src/MPC.cpp (whole file), include/qrw/MPC.hpp:19-86
//   * QPWBC   -- /root/reference/src/QPWBC.cpp:4-30,213-343,481-537, include/qrw/QPWBC.hpp:26-65
//   * InvKin  -- /root/reference/src/InvKin.cpp:3-73, include/qrw/InvKin.hpp:56-66
// with the OSQP solves replaced by the v0.6.0 restatement in osqp_restated.hpp.
//
// PARITY UNPINNED for everything that goes through OSQP (no OSQP binary/source/golden vector
// exists here -- see osqp_restated.hpp).  What IS pinned against reference code run here:
//   * the triplet->CSC value layout of ML (reference src/st_to_cc.cpp compiled into
//     oracle/_ref/libst_to_cc.so, tests/test_oracle_pins.py::test_ml_layout_matches_reference_st_to_cc)
//   * the behavioural anchors of scripts/test_mpc.py and the analytic stance forces.
//
// Matrix arguments are column-major doubles (what Eigen::MatrixXd::data() holds), one
// instance after another in the batched entry points.
#include <cstring>
#include <thread>

#include "osqp_restated.hpp"

namespace orc {

// ========================================================================================
// MPC  (src/MPC.cpp)
// ========================================================================================
struct Mpc {
  double dt, mass, mu, T_gait, h_ref;
  int n_steps, N_gait;
  double gI[3][3];
  double g8;             // g(8,0) = -9.81f * dt            (MPC.cpp:29)
  double footholds[3][4];
  double offset_CoM[3];
  std::vector<int> gait;     // N_gait x 4, row-major here
  std::vector<int> S_gait;   // 12 n
  std::vector<double> xref;  // 12 x (n+1) col-major
  double x0[12];
  double A[12][12], B[12][12];
  Csc ML, P;
  std::vector<double> Q, NK_up, NK_low;
  int i_x_B[48], i_y_B[48], i_update_B[48];
  std::vector<int> i_off;
  std::vector<double> x_f_applied;  // 24 x n col-major
  Osqp osqp;
  bool gait_overrun = false;  // reference would read out of bounds (no all-zero row in fsteps)
  int max_iter_override = 0;  // test hook (0 = OSQP default 4000)

  Mpc(double dt_in, int n_steps_in, double T_gait_in, int N_gait_in) {  // MPC.cpp:3-32
    dt = dt_in;
    n_steps = n_steps_in;
    T_gait = T_gait_in;
    N_gait = N_gait_in;
    xref.assign(12 * (1 + n_steps), 0.0);
    S_gait.assign(12 * n_steps, 0);
    x_f_applied.assign(24 * n_steps, 0.0);
    gait.assign(N_gait * 4, 0);
    mass = 2.50000279f;
    mu = 0.9f;
    offset_CoM[0] = 0.0;
    offset_CoM[1] = 0.0;
    offset_CoM[2] = -0.03;
    const double fh[12] = {0.19, 0.19, -0.19, -0.19, 0.15005, -0.15005, 0.15005, -0.15005, 0.0, 0.0, 0.0, 0.0};
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 4; ++c) footholds[r][c] = fh[4 * r + c];
    const double gi[9] = {3.09249e-2, -8.00101e-7, 1.865287e-5, -8.00101e-7, 5.106100e-2,
                          1.245813e-4, 1.865287e-5, 1.245813e-4, 6.939757e-2};
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) gI[r][c] = gi[3 * r + c];
    h_ref = 0.2027682f;
    g8 = -9.81f * dt;
    for (int r = 0; r < 12; ++r)
      for (int c = 0; c < 12; ++c) {
        A[r][c] = (r == c) ? 1.0 : 0.0;
        B[r][c] = 0.0;
      }
    for (int k = 0; k < 12; ++k) x0[k] = 0.0;
  }

  double& XR(int r, int c) { return xref[r + 12 * c]; }

  // I_inv = (R^T gI R)^-1, R = Rz(yaw)            (MPC.cpp:215-220, 425-430)
  void inertia_inverse(double yaw, double Iinv[3][3]) const {
    double c = std::cos(yaw), s = std::sin(yaw);
    double R[3][3] = {{c, -s, 0.0}, {s, c, 0.0}, {0.0, 0.0, 1.0}};
    double T[3][3], M[3][3];
    for (int i = 0; i < 3; ++i)  // T = R^T * gI
      for (int j = 0; j < 3; ++j) {
        double a = 0.0;
        for (int k = 0; k < 3; ++k) a += R[k][i] * gI[k][j];
        T[i][j] = a;
      }
    for (int i = 0; i < 3; ++i)  // M = T * R
      for (int j = 0; j < 3; ++j) {
        double a = 0.0;
        for (int k = 0; k < 3; ++k) a += T[i][k] * R[k][j];
        M[i][j] = a;
      }
    // 3x3 inverse by cofactors / determinant (what Eigen's fixed-size inverse does)
    double c00 = M[1][1] * M[2][2] - M[1][2] * M[2][1];
    double c10 = M[1][2] * M[2][0] - M[1][0] * M[2][2];
    double c20 = M[1][0] * M[2][1] - M[1][1] * M[2][0];
    double det = c00 * M[0][0] + c10 * M[0][1] + c20 * M[0][2];
    double id = 1.0 / det;
    Iinv[0][0] = c00 * id;
    Iinv[1][0] = c10 * id;
    Iinv[2][0] = c20 * id;
    Iinv[0][1] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) * id;
    Iinv[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) * id;
    Iinv[2][1] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) * id;
    Iinv[0][2] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) * id;
    Iinv[1][2] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) * id;
    Iinv[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) * id;
  }
  // B.block(9, 3i, 3, 3) = dt * (I_inv * skew(lever_i))      (MPC.cpp:224-226, 439-441)
  void fill_B_angular(const double Iinv[3][3], const double lever[3][4]) {
    for (int i = 0; i < 4; ++i) {
      double v0 = lever[0][i], v1 = lever[1][i], v2 = lever[2][i];
      double S[3][3] = {{0.0, -v2, v1}, {v2, 0.0, -v0}, {-v1, v0, 0.0}};  // getSkew, :654-658
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
          double a = 0.0;
          for (int k = 0; k < 3; ++k) a += Iinv[r][k] * S[k][c];
          B[9 + r][3 * i + c] = dt * a;
        }
    }
  }

  bool gait_row_zero(int k) const {
    if (k >= N_gait) return true;  // the reference would read past the matrix here
    for (int i = 0; i < 4; ++i)
      if (gait[4 * k + i] != 0) return false;
    return true;
  }

  void construct_gait(const double* fsteps) {  // MPC.cpp:686-701
    int k = 0;
    auto row_zero = [&](int kk) {
      for (int j = 0; j < 12; ++j)
        if (!(std::fabs(fsteps[kk + N_gait * j]) <= 1e-12)) return false;  // Eigen isZero()
      return true;
    };
    while (k < N_gait && !row_zero(k)) {
      for (int i = 0; i < 4; ++i) gait[4 * k + i] = (fsteps[k + N_gait * (3 * i)] == 0.0) ? 0 : 1;
      k++;
    }
    if (k < N_gait) {
      for (int i = 0; i < 4; ++i) gait[4 * k + i] = 0;
    } else {
      gait_overrun = true;
    }
  }
  void construct_S() {  // MPC.cpp:665-681
    int i = 0;
    while (!gait_row_zero(i)) {
      if (i < n_steps)  // the reference writes past S_gait when more rows than steps are active
        for (int b = 0; b < 4; ++b)
          for (int c = 0; c < 3; ++c) S_gait[i * 12 + 3 * b + c] = 1 - gait[4 * i + b];
      i++;
    }
  }

  void create_ML() {  // MPC.cpp:74-256
    std::vector<int> r, c;
    std::vector<double> v;
    auto add = [&](int i, int j, double val) {
      r.push_back(i);
      c.push_back(j);
      v.push_back(val);
    };
    const int n = n_steps;
    for (int k = 0; k < 12 * n; ++k) add(k, k, -1.0);
    for (int j = 0; j < 6; ++j) A[j][j + 6] = dt;
    for (int k = 0; k < n - 1; ++k) {
      for (int i = 0; i < 12; ++i) add((k + 1) * 12 + i, k * 12 + i, 1.0);
      for (int j = 0; j < 6; ++j) add((k + 1) * 12 + j, k * 12 + j + 6, dt);
    }
    double div_tmp = dt / mass;
    for (int k = 0; k < n; ++k) {
      for (int i = 0; i < 4; ++i) {
        add(12 * k + 6, 12 * (n + k) + 0 + 3 * i, div_tmp);
        add(12 * k + 7, 12 * (n + k) + 1 + 3 * i, div_tmp);
        add(12 * k + 8, 12 * (n + k) + 2 + 3 * i, div_tmp);
      }
      for (int i = 0; i < 12; ++i) {
        add(12 * k + 9, 12 * (n + k) + i, 8.0);
        add(12 * k + 10, 12 * (n + k) + i, 8.0);
        add(12 * k + 11, 12 * (n + k) + i, 8.0);
      }
    }
    for (int i = 0; i < 4; ++i) {
      B[6][0 + 3 * i] = div_tmp;
      B[7][1 + 3 * i] = div_tmp;
      B[8][2 + 3 * i] = div_tmp;
      B[9][i] = 8.0;
      B[10][i] = 8.0;
      B[11][i] = 8.0;
    }
    for (int i = 12 * n; i < 24 * n; ++i) add(i, i, 1.0);
    const int offset_L = 24 * n;
    for (int k = 0; k < n; ++k) {
      int di = offset_L + 20 * k, dj = 12 * (n + k);
      for (int i = 0; i < 4; ++i) {
        int dx = 5 * i, dy = 3 * i;
        const int a[9] = {0, 1, 2, 3, 0, 1, 2, 3, 4};
        const int b[9] = {0, 0, 1, 1, 2, 2, 2, 2, 2};
        const double cc[9] = {1.0, -1.0, 1.0, -1.0, -mu, -mu, -mu, -mu, -1};
        for (int j = 0; j < 9; ++j) add(di + dx + a[j], dj + dy + b[j], cc[j]);
      }
    }
    ML = triplet_to_csc(44 * n, 24 * n, r, c, v);

    const int i_x_tmp[12] = {6, 9, 10, 11, 7, 9, 10, 11, 8, 9, 10, 11};
    for (int k = 0; k < 4; ++k)
      for (int i = 0; i < 12; ++i) {
        i_x_B[12 * k + i] = i_x_tmp[i];
        i_y_B[12 * k + i] = (12 * k + i) / 4;
      }
    const int i_start = 30 * n - 18;
    const int i_data[12] = {0, 1, 2, 3, 7, 8, 9, 10, 14, 15, 16, 17};
    for (int k = 0; k < 4; ++k)
      for (int i = 0; i < 12; ++i) i_update_B[12 * k + i] = i_start + i_data[i] + 24 * k;

    // first fill of B: default footholds, NO CoM offset, all n steps   (MPC.cpp:213-232)
    for (int k = 0; k < n; ++k) {
      double Iinv[3][3], lever[3][4];
      inertia_inverse(XR(5, k), Iinv);
      for (int rr = 0; rr < 3; ++rr)
        for (int i = 0; i < 4; ++i) lever[rr][i] = footholds[rr][i] - XR(rr, k);
      fill_B_angular(Iinv, lever);
      int i_iter = 24 * 4 * k;
      for (int j = 0; j < 48; ++j) ML.x[i_update_B[j] + i_iter] = B[i_x_B[j]][i_y_B[j]];
    }
    construct_S();
    i_off.assign(12 * n, 0);
    i_off[0] = 4;
    for (int k = 1; k < 12 * n; ++k) i_off[k] = i_off[k - 1] + ((k - 1) % 3 == 2 ? 10 : 7);
    for (int k = 0; k < 12 * n; ++k) ML.x[i_off[k] + i_start] = S_gait[k];
  }

  // u[0:12n] = -g  (+ A*(-x0) in block 0)  + D * vec(xref[:,1:])     (MPC.cpp:261-289, 469-486)
  void fill_N() {
    const int n = n_steps;
    NK_up.assign(44 * n, 0.0);
    for (int k = 0; k < n; ++k) NK_up[12 * k + 8] = -g8;
    for (int rr = 0; rr < 12; ++rr) {
      double a = 0.0;
      for (int cc = 0; cc < 12; ++cc) a += A[rr][cc] * (-x0[cc]);
      NK_up[rr] += a;
    }
    // D = I with -A sub-diagonal blocks; dense product in the reference, same non-zero terms
    // in the same (column) order here
    for (int k = 0; k < n; ++k)
      for (int i = 0; i < 12; ++i) {
        double s = 0.0;
        if (k > 0) {
          s += -1.0 * XR(i, k);  // column 12(k-1)+i of D times xref_col = xref(:,k)
          if (i < 6) s += -dt * XR(i + 6, k);
        }
        s += 1.0 * XR(i, k + 1);
        NK_up[12 * k + i] += s;
      }
  }
  void create_NK() {  // MPC.cpp:261-312
    const int n = n_steps;
    fill_N();
    NK_low.assign(44 * n, 0.0);
    for (int i = 0; i < 24 * n; ++i) NK_low[i] = NK_up[i];
    for (int i = 0; i < 20 * n; ++i)
      NK_low[24 * n + i] = -std::numeric_limits<double>::infinity();
    for (int k = 0; (4 + 5 * k) < 20 * n; ++k) NK_low[24 * n + 4 + 5 * k] = -25.0;
  }
  void update_NK() {  // MPC.cpp:469-496
    fill_N();
    for (int i = 0; i < 24 * n_steps; ++i) NK_low[i] = NK_up[i];
  }
  void create_weight_matrices() {  // MPC.cpp:317-391
    std::vector<int> r, c;
    std::vector<double> v;
    const double w[12] = {2.0f, 2.0f, 20.0f, 0.25f, 0.25f, 10.0f, 0.2f, 0.2f, 0.2f, 0.0f, 0.0f, 0.3f};
    const int n = n_steps;
    for (int k = 0; k < n; ++k)
      for (int i = 0; i < 12; ++i) {
        r.push_back(12 * k + i);
        c.push_back(12 * k + i);
        v.push_back(w[i]);
      }
    for (int k = n; k < 2 * n; ++k)
      for (int i = 0; i < 12; ++i) {
        r.push_back(12 * k + i);
        c.push_back(12 * k + i);
        v.push_back(5e-5f);
      }
    P = triplet_to_csc(24 * n, 24 * n, r, c, v);
    Q.assign(24 * n, 0.0);
  }
  void update_ML(const double* fsteps) {  // MPC.cpp:418-464
    int j = 0;
    while (!gait_row_zero(j)) {
      int k = j;
      if (k < n_steps) {
        double Iinv[3][3], lever[3][4];
        inertia_inverse(XR(5, k), Iinv);
        for (int rr = 0; rr < 3; ++rr)
          for (int i = 0; i < 4; ++i)
            lever[rr][i] = fsteps[j + N_gait * (3 * i + rr)] - (XR(rr, k) + offset_CoM[rr]);
        fill_B_angular(Iinv, lever);
        int i_iter = 24 * 4 * k;
        for (int i = 0; i < 48; ++i) ML.x[i_update_B[i] + i_iter] = B[i_x_B[i]][i_y_B[i]];
      }
      j++;
    }
    construct_S();
    const int i_start = 30 * n_steps - 18;
    for (int k = 0; k < 12 * n_steps; ++k) ML.x[i_off[k] + i_start] = S_gait[k];
  }

  Settings settings() const {  // MPC.cpp:527-539 on top of osqp_set_default_settings
    Settings s;
    s.sigma = 1e-6;
    s.eps_abs = 1e-6;
    s.eps_rel = 1e-6;
    s.eps_prim_inf = 1e-5;
    s.eps_dual_inf = 1e-4;
    s.alpha = 1.6;
    s.adaptive_rho = 1;
    s.adaptive_rho_interval = 200;
    s.adaptive_rho_tolerance = 5.0;
    if (max_iter_override > 0) s.max_iter = max_iter_override;
    return s;
  }

  int run(int num_iter, const double* xref_in, const double* fsteps_in) {  // MPC.cpp:626-649
    construct_gait(fsteps_in);
    std::memcpy(xref.data(), xref_in, sizeof(double) * 12 * (n_steps + 1));
    for (int k = 0; k < 12; ++k) x0[k] = xref_in[k];
    if (num_iter == 0) {
      create_ML();
      create_NK();
      create_weight_matrices();
      osqp.setup(P, ML, Q.data(), NK_low.data(), NK_up.data(), settings());
    } else {
      update_ML(fsteps_in);
      update_NK();
      osqp.update_A(ML.x.data());
      osqp.update_bounds(NK_low.data(), NK_up.data());
    }
    osqp.solve();
    // retrieve_result, MPC.cpp:569-599
    for (int i = 0; i < n_steps; ++i)
      for (int k = 0; k < 12; ++k) {
        x_f_applied[k + 24 * i] = osqp.sol_x[k + 12 * i] + XR(k, 1 + i);
        x_f_applied[k + 12 + 24 * i] = osqp.sol_x[12 * (n_steps + i) + k];
      }
    return 0;
  }
};

// ========================================================================================
// pseudoInverse (include/qrw/InvKin.hpp:60-66): V * diag(sv > tol ? 1/sv : 0) * U^T with
// tol = eps * max(rows, cols) * sv_max, from a Jacobi SVD.  One-sided (Hestenes) Jacobi here.
// Column-major N x N in, column-major out.
// ========================================================================================
template <int N>
void pseudo_inverse(const double* a, double* out) {
  double U[N][N], V[N][N];  // U[col][row]
  for (int c = 0; c < N; ++c)
    for (int r = 0; r < N; ++r) {
      U[c][r] = a[r + N * c];
      V[c][r] = (r == c) ? 1.0 : 0.0;
    }
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < N - 1; ++p)
      for (int q = p + 1; q < N; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int r = 0; r < N; ++r) {
          alpha += U[p][r] * U[p][r];
          beta += U[q][r] * U[q][r];
          gamma += U[p][r] * U[q][r];
        }
        if (gamma == 0.0 || std::fabs(gamma) <= 1e-15 * std::sqrt(alpha * beta)) continue;
        rotated = true;
        double zeta = (beta - alpha) / (2.0 * gamma);
        double t = ((zeta >= 0) ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
        double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
        for (int r = 0; r < N; ++r) {
          double up = U[p][r], uq = U[q][r];
          U[p][r] = cs * up - sn * uq;
          U[q][r] = sn * up + cs * uq;
          double vp = V[p][r], vq = V[q][r];
          V[p][r] = cs * vp - sn * vq;
          V[q][r] = sn * vp + cs * vq;
        }
      }
    if (!rotated) break;
  }
  double sv[N], svmax = 0.0;
  for (int c = 0; c < N; ++c) {
    double s = 0;
    for (int r = 0; r < N; ++r) s += U[c][r] * U[c][r];
    sv[c] = std::sqrt(s);
    svmax = std::max(svmax, sv[c]);
  }
  double tol = std::numeric_limits<double>::epsilon() * (double)N * svmax;
  for (int i = 0; i < N * N; ++i) out[i] = 0.0;
  for (int c = 0; c < N; ++c) {
    if (!(sv[c] > tol)) continue;
    double inv2 = 1.0 / (sv[c] * sv[c]);  // V[:,c] * (1/sv) * (U[:,c]/sv)^T
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < N; ++j) out[i + N * j] += V[c][i] * inv2 * U[c][j];
  }
}

// ========================================================================================
// QPWBC  (src/QPWBC.cpp)
// ========================================================================================
struct Wbc {
  bool initialized = false;
  double G[20][12];
  double Y[36], Yinv[36], X[6][12], A[6][12], gamma[6], H[12][12], g[12];
  double f_res[12], ddq_res[6];
  Csc ML, P;
  double Q[12], v_NK_up[20], v_NK_low[20];
  Osqp osqp;

  Wbc() {  // QPWBC.cpp:4-30
    const double mu = 0.9;
    double SC[5][3] = {{0}};
    const int a[9] = {0, 1, 2, 3, 0, 1, 2, 3, 4};
    const int b[9] = {0, 0, 1, 1, 2, 2, 2, 2, 2};
    const double c[9] = {1.0, -1.0, 1.0, -1.0, -mu, -mu, -mu, -mu, -1};
    for (int i = 0; i <= 8; ++i) SC[a[i]][b[i]] = -c[i];
    std::memset(G, 0, sizeof(G));
    for (int i = 0; i < 4; ++i)
      for (int r = 0; r < 5; ++r)
        for (int cc = 0; cc < 3; ++cc) G[5 * i + r][3 * i + cc] = SC[r][cc];
    for (int i = 0; i < 20; ++i) {
      v_NK_up[i] = 25.0;
      v_NK_low[i] = 0.0;
    }
    std::memset(f_res, 0, sizeof(f_res));
    std::memset(ddq_res, 0, sizeof(ddq_res));
    std::memset(H, 0, sizeof(H));
  }
  void create_matrices() {  // QPWBC.cpp:32-211
    std::vector<int> r, c;
    std::vector<double> v;
    for (int i = 0; i < 20; ++i)
      for (int j = 0; j < 12; ++j) {
        r.push_back(i);
        c.push_back(j);
        v.push_back(G[i][j]);
      }
    ML = triplet_to_csc(20, 12, r, c, v);
    r.clear();
    c.clear();
    v.clear();
    for (int i = 0; i < 12; ++i)
      for (int j = i; j < 12; ++j) {
        r.push_back(i);
        c.push_back(j);
        v.push_back(1.0);
      }
    P = triplet_to_csc(12, 12, r, c, v);
    for (int i = 0; i < 12; ++i) Q[i] = 0.0;
  }
  // M 18x18, Jc 12x18 column-major                                  (QPWBC.cpp:481-498)
  void compute_matrices(const double* M, const double* Jc, const double* f_cmd, const double* RNEA) {
    for (int c = 0; c < 6; ++c)
      for (int r = 0; r < 6; ++r) Y[r + 6 * c] = M[r + 18 * c];
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 12; ++c) X[r][c] = Jc[c + 12 * r];
    pseudo_inverse<6>(Y, Yinv);
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 12; ++c) {
        double s = 0;
        for (int k = 0; k < 6; ++k) s += Yinv[r + 6 * k] * X[k][c];
        A[r][c] = s;
      }
    double t[6];
    for (int r = 0; r < 6; ++r) {
      double s = 0;
      for (int c = 0; c < 12; ++c) s += X[r][c] * f_cmd[c];
      t[r] = s - RNEA[r];
    }
    for (int r = 0; r < 6; ++r) {
      double s = 0;
      for (int k = 0; k < 6; ++k) s += Yinv[r + 6 * k] * t[k];
      gamma[r] = s;
    }
    // H = A^T Q1 A + Q2,  g = A^T Q1 gamma,  Q1 = 0.1 I6, Q2 = 5 I12   (QPWBC.hpp:26-27)
    for (int i = 0; i < 12; ++i) {
      for (int j = 0; j < 12; ++j) {
        double s = 0;
        for (int k = 0; k < 6; ++k) s += (A[k][i] * 0.1) * A[k][j];
        H[i][j] = s + (i == j ? 5.0 : 0.0);
      }
      double s = 0;
      for (int k = 0; k < 6; ++k) s += (A[k][i] * 0.1) * gamma[k];
      g[i] = s;
    }
  }
  Settings settings() const {  // QPWBC.cpp:239-251
    Settings s;
    s.eps_abs = (float)1e-5;
    s.eps_rel = (float)1e-5;
    s.adaptive_rho = 1;
    s.adaptive_rho_interval = 200;
    s.adaptive_rho_tolerance = (float)5.0;
    return s;
  }
  int run(const double* M, const double* Jc, const double* f_cmd, const double* RNEA) {  // :310-343
    if (!initialized) create_matrices();
    compute_matrices(M, Jc, f_cmd, RNEA);
    int cpt = 0;  // update_PQ, :520-537
    for (int i = 0; i < 12; ++i)
      for (int j = 0; j <= i; ++j) P.x[cpt++] = H[j][i];
    for (int i = 0; i < 12; ++i) Q[i] = g[i];
    for (int i = 0; i < 20; ++i) {
      double Gf = 0;
      for (int j = 0; j < 12; ++j) Gf += G[i][j] * f_cmd[j];
      v_NK_low[i] = -Gf;
      v_NK_up[i] = -Gf + 25.0;
    }
    if (!initialized) {  // call_solver, :213-275
      osqp.setup(P, ML, Q, v_NK_low, v_NK_up, settings());
      initialized = true;
    } else {
      osqp.update_P(P.x.data());
      osqp.update_lin_cost(Q);
      osqp.update_upper_bound(v_NK_up);
      osqp.update_lower_bound(v_NK_low);
    }
    osqp.solve();
    for (int k = 0; k < 12; ++k) f_res[k] = osqp.sol_x[k];  // retrieve_result, :277-297
    for (int r = 0; r < 6; ++r) {
      double s = 0;
      for (int c = 0; c < 12; ++c) s += A[r][c] * f_res[c];
      ddq_res[r] = s + gamma[r];
    }
    for (int k = 0; k < 12; ++k) f_res[k] += f_cmd[k];
    return 0;
  }
};

// ========================================================================================
// InvKin  (src/InvKin.cpp:23-73)
// ========================================================================================
struct InvKinO {
  double ddq[12], q_step[12], dq_cmd[12];
  const double Kp = 100.0;
  const double Kd = 2.0 * std::sqrt(100.0);
  InvKinO() {
    std::memset(ddq, 0, sizeof(ddq));
    std::memset(q_step, 0, sizeof(q_step));
    std::memset(dq_cmd, 0, sizeof(dq_cmd));
  }
  // contacts 1x4; goals/vgoals/agoals 3x4; posf/vf/wf/af 4x3; Jf 12x12 -- all column-major
  void run(const double* contacts, const double* goals, const double* vgoals, const double* agoals,
           const double* posf, const double* vf, const double* wf, const double* af, const double* Jf) {
    double acc[12], x_err[12], dx_r[12];
    for (int i = 0; i < 4; ++i) {
      double e[3], a[3], v[3], w[3];
      for (int c = 0; c < 3; ++c) {
        v[c] = vf[i + 4 * c];
        w[c] = wf[i + 4 * c];
        e[c] = goals[c + 3 * i] - posf[i + 4 * c];
        a[c] = Kp * e[c] - Kd * (v[c] - vgoals[c + 3 * i]) + agoals[c + 3 * i];
      }
      if (contacts[i] != 0.0)
        for (int c = 0; c < 3; ++c) a[c] *= 0.0;
      double cr[3] = {w[1] * v[2] - w[2] * v[1], w[2] * v[0] - w[0] * v[2], w[0] * v[1] - w[1] * v[0]};
      for (int c = 0; c < 3; ++c) {
        a[c] -= af[i + 4 * c] + cr[c];
        acc[3 * i + c] = a[c];
        x_err[3 * i + c] = e[c];
        dx_r[3 * i + c] = vgoals[c + 3 * i];
      }
    }
    for (int i = 0; i < 4; ++i) {
      double M[3][3], I[3][3];
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) M[r][c] = Jf[(3 * i + r) + 12 * (3 * i + c)];
      double c00 = M[1][1] * M[2][2] - M[1][2] * M[2][1];
      double c10 = M[1][2] * M[2][0] - M[1][0] * M[2][2];
      double c20 = M[1][0] * M[2][1] - M[1][1] * M[2][0];
      double det = c00 * M[0][0] + c10 * M[0][1] + c20 * M[0][2];
      double id = 1.0 / det;
      I[0][0] = c00 * id;
      I[1][0] = c10 * id;
      I[2][0] = c20 * id;
      I[0][1] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) * id;
      I[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) * id;
      I[2][1] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) * id;
      I[0][2] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) * id;
      I[1][2] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) * id;
      I[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) * id;
      for (int r = 0; r < 3; ++r) {
        double s0 = 0, s1 = 0, s2 = 0;
        for (int c = 0; c < 3; ++c) {
          s0 += I[r][c] * acc[3 * i + c];
          s1 += I[r][c] * dx_r[3 * i + c];
          s2 += I[r][c] * x_err[3 * i + c];
        }
        ddq[3 * i + r] = s0;
        dq_cmd[3 * i + r] = s1;
        q_step[3 * i + r] = s2;
      }
    }
  }
};

}  // namespace orc

// ========================================================================================
// C API (ctypes) -- single instance and instance-major batches.
// ========================================================================================
using namespace orc;

struct MpcBatch {
  std::vector<Mpc*> inst;
  int n_steps, N_gait;
};
struct WbcBatch {
  std::vector<Wbc*> inst;
};

template <class F>
static void parallel_for(int B, int nthreads, F f) {
  if (nthreads <= 1 || B <= 1) {
    for (int b = 0; b < B; ++b) f(b);
    return;
  }
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t)
    th.emplace_back([=]() {
      for (int b = t; b < B; b += nthreads) f(b);
    });
  for (auto& t : th) t.join();
}

extern "C" {

void* orc_mpc_create(int batch, double dt, int n_steps, double T_gait, int N_gait) {
  auto* h = new MpcBatch();
  h->n_steps = n_steps;
  h->N_gait = N_gait;
  for (int b = 0; b < batch; ++b) h->inst.push_back(new Mpc(dt, n_steps, T_gait, N_gait));
  return h;
}
void orc_mpc_destroy(void* hv) {
  auto* h = (MpcBatch*)hv;
  for (auto* p : h->inst) delete p;
  delete h;
}
// xref: [B][12*(n+1)] col-major per instance; fsteps: [B][N_gait*12] col-major per instance
int orc_mpc_run(void* hv, int num_iter, const double* xref, const double* fsteps, int nthreads) {
  auto* h = (MpcBatch*)hv;
  const int sx = 12 * (h->n_steps + 1), sf = 12 * h->N_gait;
  parallel_for((int)h->inst.size(), nthreads,
               [=](int b) { h->inst[b]->run(num_iter, xref + (size_t)b * sx, fsteps + (size_t)b * sf); });
  return 0;
}
void orc_mpc_result(void* hv, double* out) {  // [B][24*n] col-major per instance
  auto* h = (MpcBatch*)hv;
  const int s = 24 * h->n_steps;
  for (size_t b = 0; b < h->inst.size(); ++b)
    std::memcpy(out + b * s, h->inst[b]->x_f_applied.data(), sizeof(double) * s);
}
void orc_mpc_gait(void* hv, double* out) {  // [B][N_gait*4] col-major per instance (get_gait)
  auto* h = (MpcBatch*)hv;
  const int N = h->N_gait;
  for (size_t b = 0; b < h->inst.size(); ++b)
    for (int k = 0; k < N; ++k)
      for (int i = 0; i < 4; ++i) out[b * 4 * N + k + N * i] = (double)h->inst[b]->gait[4 * k + i];
}
void orc_mpc_sgait(void* hv, double* out) {  // [B][12 n]
  auto* h = (MpcBatch*)hv;
  const int s = 12 * h->n_steps;
  for (size_t b = 0; b < h->inst.size(); ++b)
    for (int k = 0; k < s; ++k) out[b * s + k] = (double)h->inst[b]->S_gait[k];
}
// per-instance solver info: iters[B], status[B], rho[B], n_factor[B]
void orc_mpc_info(void* hv, int* iters, int* status, double* rho, int* nfactor) {
  auto* h = (MpcBatch*)hv;
  for (size_t b = 0; b < h->inst.size(); ++b) {
    iters[b] = h->inst[b]->osqp.iter;
    status[b] = h->inst[b]->osqp.status;
    rho[b] = h->inst[b]->osqp.st.rho;
    nfactor[b] = h->inst[b]->osqp.n_factor;
  }
}
// Introspection of ONE instance (layout / assembly / scaling tests).
int orc_mpc_nnz(void* hv, int b) { return ((MpcBatch*)hv)->inst[b]->ML.nnz(); }
void orc_mpc_ml(void* hv, int b, double* x, int* rowidx, int* colptr) {
  auto* m = ((MpcBatch*)hv)->inst[b];
  if (x) std::memcpy(x, m->ML.x.data(), sizeof(double) * m->ML.nnz());
  if (rowidx) std::memcpy(rowidx, m->ML.i.data(), sizeof(int) * m->ML.nnz());
  if (colptr) std::memcpy(colptr, m->ML.p.data(), sizeof(int) * (m->ML.n + 1));
}
void orc_mpc_bounds(void* hv, int b, double* l, double* u) {
  auto* m = ((MpcBatch*)hv)->inst[b];
  std::memcpy(l, m->NK_low.data(), sizeof(double) * m->NK_low.size());
  std::memcpy(u, m->NK_up.data(), sizeof(double) * m->NK_up.size());
}
// scaled problem data + iterates held by the solver after the last call
void orc_mpc_solver_state(void* hv, int b, double* D, double* E, double* c, double* x, double* z,
                          double* y, double* As, double* Pdiag_s, double* ls, double* us) {
  auto& o = ((MpcBatch*)hv)->inst[b]->osqp;
  if (D) std::memcpy(D, o.D.data(), sizeof(double) * o.n);
  if (E) std::memcpy(E, o.E.data(), sizeof(double) * o.m);
  if (c) *c = o.c;
  if (x) std::memcpy(x, o.x.data(), sizeof(double) * o.n);
  if (z) std::memcpy(z, o.z.data(), sizeof(double) * o.m);
  if (y) std::memcpy(y, o.y.data(), sizeof(double) * o.m);
  if (As) std::memcpy(As, o.A.x.data(), sizeof(double) * o.A.nnz());
  if (Pdiag_s) std::memcpy(Pdiag_s, o.P.x.data(), sizeof(double) * o.P.nnz());
  if (ls) std::memcpy(ls, o.l.data(), sizeof(double) * o.m);
  if (us) std::memcpy(us, o.u.data(), sizeof(double) * o.m);
}
void orc_mpc_set_trace(void* hv, int b, int on) { ((MpcBatch*)hv)->inst[b]->osqp.keep_trace = on != 0; }
int orc_mpc_trace(void* hv, int b, double* out, int cap) {  // rows of (iter, pri, dua, rho)
  auto& t = ((MpcBatch*)hv)->inst[b]->osqp.trace;
  int nrow = (int)t.size() / 4;
  if (out)
    for (int i = 0; i < std::min(cap, nrow) * 4; ++i) out[i] = t[i];
  return nrow;
}
void orc_mpc_set_max_iter(void* hv, int max_iter) {  // test hook only
  for (auto* p : ((MpcBatch*)hv)->inst) {
    p->max_iter_override = max_iter;
    p->osqp.st.max_iter = max_iter;
  }
}

// ---- QPWBC ----
void* orc_wbc_create(int batch) {
  auto* h = new WbcBatch();
  for (int b = 0; b < batch; ++b) h->inst.push_back(new Wbc());
  return h;
}
void orc_wbc_destroy(void* hv) {
  auto* h = (WbcBatch*)hv;
  for (auto* p : h->inst) delete p;
  delete h;
}
// M [B][18*18], Jc [B][12*18], f_cmd [B][12], RNEA [B][6]; all column-major per instance
int orc_wbc_run(void* hv, const double* M, const double* Jc, const double* f_cmd, const double* RNEA,
                int nthreads) {
  auto* h = (WbcBatch*)hv;
  parallel_for((int)h->inst.size(), nthreads, [=](int b) {
    h->inst[b]->run(M + (size_t)b * 324, Jc + (size_t)b * 216, f_cmd + (size_t)b * 12, RNEA + (size_t)b * 6);
  });
  return 0;
}
void orc_wbc_result(void* hv, double* f_res, double* ddq_res, double* H) {  // [B][12],[B][6],[B][144] col-major
  auto* h = (WbcBatch*)hv;
  for (size_t b = 0; b < h->inst.size(); ++b) {
    auto* w = h->inst[b];
    if (f_res) std::memcpy(f_res + 12 * b, w->f_res, sizeof(double) * 12);
    if (ddq_res) std::memcpy(ddq_res + 6 * b, w->ddq_res, sizeof(double) * 6);
    if (H)
      for (int i = 0; i < 12; ++i)
        for (int j = 0; j < 12; ++j) H[144 * b + i + 12 * j] = w->H[i][j];
  }
}
void orc_wbc_info(void* hv, int* iters, int* status, double* rho) {
  auto* h = (WbcBatch*)hv;
  for (size_t b = 0; b < h->inst.size(); ++b) {
    iters[b] = h->inst[b]->osqp.iter;
    status[b] = h->inst[b]->osqp.status;
    rho[b] = h->inst[b]->osqp.st.rho;
  }
}
void orc_pinv6(const double* a, double* out) { pseudo_inverse<6>(a, out); }

// ---- InvKin (stateless apart from its outputs) ----
// batch of B; every array instance-major, column-major inside an instance.
void orc_invkin_run(int B, const double* contacts, const double* goals, const double* vgoals,
                    const double* agoals, const double* posf, const double* vf, const double* wf,
                    const double* af, const double* Jf, double* ddq, double* dq_cmd, double* q_step) {
  for (int b = 0; b < B; ++b) {
    InvKinO k;
    k.run(contacts + 4 * b, goals + 12 * b, vgoals + 12 * b, agoals + 12 * b, posf + 12 * b, vf + 12 * b,
          wf + 12 * b, af + 12 * b, Jf + 144 * b);
    std::memcpy(ddq + 12 * b, k.ddq, sizeof(k.ddq));
    std::memcpy(dq_cmd + 12 * b, k.dq_cmd, sizeof(k.dq_cmd));
    std::memcpy(q_step + 12 * b, k.q_step, sizeof(k.q_step));
  }
}

// ---- generic OSQP-restated entry point (for the independent-solver tests) ----
// dense column-major P (n x n, symmetric) and A (m x n); zeros are dropped from the pattern.
int orc_osqp_dense(int n, int m, const double* Pd, const double* Ad, const double* q, const double* l,
                   const double* u, double eps, int max_iter, double* x_out, double* y_out, int* iters) {
  std::vector<int> r, c;
  std::vector<double> v;
  for (int j = 0; j < n; ++j)
    for (int i = 0; i <= j; ++i)
      if (Pd[i + n * j] != 0.0 || i == j) {
        r.push_back(i);
        c.push_back(j);
        v.push_back(Pd[i + n * j]);
      }
  Csc P = triplet_to_csc(n, n, r, c, v);
  r.clear();
  c.clear();
  v.clear();
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < m; ++i)
      if (Ad[i + m * j] != 0.0) {
        r.push_back(i);
        c.push_back(j);
        v.push_back(Ad[i + m * j]);
      }
  Csc A = triplet_to_csc(m, n, r, c, v);
  Settings s;
  s.eps_abs = eps;
  s.eps_rel = eps;
  s.max_iter = max_iter;
  s.adaptive_rho_interval = 200;
  Osqp o;
  o.setup(P, A, q, l, u, s);
  o.solve();
  std::memcpy(x_out, o.sol_x.data(), sizeof(double) * n);
  std::memcpy(y_out, o.sol_y.data(), sizeof(double) * m);
  *iters = o.iter;
  return o.status;
}

// The same with alpha and the adaptive-rho interval chosen by the caller (0 = OSQP's rule without PROFILING: 4 x
// check_termination), OSQP's default tolerances unless eps says otherwise, and what OSQP prints returned: the last iterate
// (also when max_iter stops the run -- max_iter = 1 gives the figures of OSQP's "iter 1" line), info5 = {pri_res, dua_res, rho,
// number of rho updates, rho estimate at the end}.  tests/test_oracle_pins.py compares them with the output printed in OSQP's
// own documentation.
int orc_osqp_dense_ex(int n, int m, const double* Pd, const double* Ad, const double* q, const double* l, const double* u,
                      double eps, double alpha, int max_iter, int adaptive_rho_interval, double* x_out, double* y_out,
                      int* iters, double* info5) {
  std::vector<int> r, c;
  std::vector<double> v;
  for (int j = 0; j < n; ++j)
    for (int i = 0; i <= j; ++i)
      if (Pd[i + n * j] != 0.0 || i == j) r.push_back(i), c.push_back(j), v.push_back(Pd[i + n * j]);
  Csc P = triplet_to_csc(n, n, r, c, v);
  r.clear(), c.clear(), v.clear();
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < m; ++i)
      if (Ad[i + m * j] != 0.0) r.push_back(i), c.push_back(j), v.push_back(Ad[i + m * j]);
  Csc A = triplet_to_csc(m, n, r, c, v);
  Settings s;
  s.eps_abs = eps;
  s.eps_rel = eps;
  s.alpha = alpha;
  s.max_iter = max_iter;
  s.adaptive_rho_interval = adaptive_rho_interval;
  Osqp o;
  o.setup(P, A, q, l, u, s);
  o.solve();
  for (int j = 0; j < n; ++j) x_out[j] = o.st.scaling ? o.D[j] * o.x[j] : o.x[j];
  for (int i = 0; i < m; ++i) y_out[i] = o.st.scaling ? o.E[i] * o.y[i] * o.cinv : o.y[i];
  *iters = o.iter;
  info5[0] = o.pri_res, info5[1] = o.dua_res, info5[2] = o.st.rho, info5[3] = (double)o.rho_updates;
  info5[4] = o.compute_rho_estimate();
  return o.status;
}

}  // extern "C"
