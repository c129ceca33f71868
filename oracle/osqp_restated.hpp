// oracle/osqp_restated.hpp
//
// TEST INFRASTRUCTURE ONLY -- CPU oracle.  Nothing under oracle/ is part of the shipped
// product path; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may build, load or call it.
//
// What this file restates
// -----------------------
// The reference's two QPs (src/MPC.cpp:501-564, src/QPWBC.cpp:213-275) are solved by OSQP,
// a third-party dependency that is NOT vendored in /root/reference and NOT pinned
// (README.md:21-31 clones upstream master; the API used -- osqp_setup(&work,data,settings),
// osqp_update_A/P/bounds -- fixes the range to v0.6.0..v0.6.3).  This header restates the
// published OSQP v0.6.0 algorithm (Stellato et al., "OSQP: an operator splitting solver for
// quadratic programs", Math. Prog. Comp. 2020, and the v0.6.0 C sources: osqp.c, auxil.c,
// scaling.c, lin_alg.c, kkt.c, qdldl_interface.c, qdldl.c) in the build configuration the
// reference README prescribes: PRINTING OFF, PROFILING OFF, DLONG OFF, EMBEDDED off,
// c_float = double.  The fill-reducing ordering is a plain minimum-degree ordering instead of
// SuiteSparse AMD: any symmetric permutation of a quasi-definite KKT matrix gives the same
// solution up to round-off.
//
// PARITY: there is no OSQP binary, no OSQP source and no reference-held golden vector for these solves in this
// environment, so the restatement cannot be run against OSQP itself (the MPC / QPWBC solves stay "parity unpinned" in that
// sense).  What it IS checked against:
//   * the ONE OSQP output available: the run printed in OSQP's own documentation (Examples -> Setup and solve, v0.6.0,
//     alpha = 1.0).  The restatement reproduces every printed digit of its "iter 1" line, its final line (50 iterations,
//     rho 0.1 -> 1.38), the solution and the "optimal rho estimate" (tests/test_oracle_pins.py::
//     test_restated_osqp_reproduces_the_run_printed_in_osqps_documentation; the figures come from the builder's memory of
//     that page -- no network here -- and were not fitted);
//   * a second, dense-numpy implementation of the iteration written independently (tests/test_osqp_second_implementation.py:
//     identical ADMM iteration counts on random QPs with equality / one-sided rows);
//   * independent exact QP solutions (tests/test_oracle_pins.py: KKT conditions of random QPs; tests/test_exact_qp.py: an
//     interior-point solve of the assembled MPC QP) and the analytic anchors of SURVEY.md section 4.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <limits>
#include <set>
#include <vector>

namespace orc {

// ----------------------------------------------------------------------------------------
// Compressed-sparse-column matrix and triplet conversion.
// Triplet -> CSC follows the semantics of the reference's vendored converter
// (src/st_to_cc.cpp:1622 st_to_cc_size, :1675 st_to_cc_index, :1763 st_to_cc_values):
// entries sorted by (column, row), duplicates summed.  tests/test_oracle_pins.py
// (test_ml_layout_matches_reference_st_to_cc) checks this against the reference converter compiled into oracle/_ref/.
// ----------------------------------------------------------------------------------------
struct Csc {
  int m = 0, n = 0;
  std::vector<int> p, i;
  std::vector<double> x;
  int nnz() const { return p.empty() ? 0 : p[n]; }
};

inline Csc triplet_to_csc(int m, int n, const std::vector<int>& r, const std::vector<int>& c,
                          const std::vector<double>& v) {
  const int nst = (int)r.size();
  std::vector<int> ord(nst);
  for (int k = 0; k < nst; ++k) ord[k] = k;
  std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) {
    if (c[a] != c[b]) return c[a] < c[b];
    return r[a] < r[b];
  });
  Csc M;
  M.m = m;
  M.n = n;
  M.p.assign(n + 1, 0);
  int last_r = -1, last_c = -1;
  for (int k = 0; k < nst; ++k) {
    int t = ord[k];
    if (r[t] == last_r && c[t] == last_c) {
      M.x.back() += v[t];
    } else {
      M.i.push_back(r[t]);
      M.x.push_back(v[t]);
      M.p[c[t] + 1]++;
      last_r = r[t];
      last_c = c[t];
    }
  }
  for (int j = 0; j < n; ++j) M.p[j + 1] += M.p[j];
  return M;
}

// ----------------------------------------------------------------------------------------
// lin_alg.c equivalents (same loop order as OSQP v0.6.0 so round-off follows it).
// ----------------------------------------------------------------------------------------
inline void mat_vec(const Csc& A, const double* x, double* y, int plus_eq) {
  if (!plus_eq)
    for (int i = 0; i < A.m; ++i) y[i] = 0.0;
  if (A.nnz() == 0) return;
  if (plus_eq == -1) {
    for (int j = 0; j < A.n; ++j)
      for (int k = A.p[j]; k < A.p[j + 1]; ++k) y[A.i[k]] -= A.x[k] * x[j];
  } else {
    for (int j = 0; j < A.n; ++j)
      for (int k = A.p[j]; k < A.p[j + 1]; ++k) y[A.i[k]] += A.x[k] * x[j];
  }
}
inline void mat_tpose_vec(const Csc& A, const double* x, double* y, int plus_eq, int skip_diag) {
  if (!plus_eq)
    for (int j = 0; j < A.n; ++j) y[j] = 0.0;
  if (A.nnz() == 0) return;
  const double sgn = (plus_eq == -1) ? -1.0 : 1.0;
  for (int j = 0; j < A.n; ++j)
    for (int k = A.p[j]; k < A.p[j + 1]; ++k) {
      int i = A.i[k];
      if (skip_diag && i == j) continue;
      if (sgn < 0)
        y[j] -= A.x[k] * x[i];
      else
        y[j] += A.x[k] * x[i];
    }
}
inline double vec_norm_inf(const double* v, int l) {
  double mx = 0.0;
  for (int i = 0; i < l; ++i) {
    double a = std::fabs(v[i]);
    if (a > mx) mx = a;
  }
  return mx;
}
inline double vec_scaled_norm_inf(const double* S, const double* v, int l) {
  double mx = 0.0;
  for (int i = 0; i < l; ++i) {
    double a = std::fabs(S[i] * v[i]);
    if (a > mx) mx = a;
  }
  return mx;
}
inline void mat_inf_norm_cols(const Csc& M, double* E) {
  for (int j = 0; j < M.n; ++j) E[j] = 0.0;
  for (int j = 0; j < M.n; ++j)
    for (int k = M.p[j]; k < M.p[j + 1]; ++k) E[j] = std::max(std::fabs(M.x[k]), E[j]);
}
inline void mat_inf_norm_rows(const Csc& M, double* E) {
  for (int i = 0; i < M.m; ++i) E[i] = 0.0;
  for (int j = 0; j < M.n; ++j)
    for (int k = M.p[j]; k < M.p[j + 1]; ++k) E[M.i[k]] = std::max(std::fabs(M.x[k]), E[M.i[k]]);
}
inline void mat_inf_norm_cols_sym_triu(const Csc& M, double* E) {
  for (int j = 0; j < M.n; ++j) E[j] = 0.0;
  for (int j = 0; j < M.n; ++j)
    for (int k = M.p[j]; k < M.p[j + 1]; ++k) {
      int i = M.i[k];
      double a = std::fabs(M.x[k]);
      E[j] = std::max(a, E[j]);
      if (i != j) E[i] = std::max(a, E[i]);
    }
}
inline void mat_premult_diag(Csc& A, const double* d) {
  for (int j = 0; j < A.n; ++j)
    for (int k = A.p[j]; k < A.p[j + 1]; ++k) A.x[k] *= d[A.i[k]];
}
inline void mat_postmult_diag(Csc& A, const double* d) {
  for (int j = 0; j < A.n; ++j)
    for (int k = A.p[j]; k < A.p[j + 1]; ++k) A.x[k] *= d[j];
}

// ----------------------------------------------------------------------------------------
// Sparse LDL^T of a symmetric quasi-definite matrix given by its upper triangle (CSC),
// up-looking with an elimination tree -- the QDLDL algorithm (qdldl.c: QDLDL_etree,
// QDLDL_factor, QDLDL_solve) restated.
// ----------------------------------------------------------------------------------------
struct Ldl {
  int n = 0;
  std::vector<int> etree, Lnz, Lp, Li;
  std::vector<double> Lx, D, Dinv;
  // scratch
  std::vector<int> mark, ylist, ebuf, nextcol;
  std::vector<double> yv;

  bool symbolic(const Csc& K) {
    n = K.n;
    etree.assign(n, -1);
    Lnz.assign(n, 0);
    std::vector<int> work(n, 0);
    for (int j = 0; j < n; ++j) {
      if (K.p[j] == K.p[j + 1]) return false;
      work[j] = j;
      for (int q = K.p[j]; q < K.p[j + 1]; ++q) {
        int i = K.i[q];
        if (i > j) return false;
        while (work[i] != j) {
          if (etree[i] == -1) etree[i] = j;
          Lnz[i]++;
          work[i] = j;
          i = etree[i];
        }
      }
    }
    Lp.assign(n + 1, 0);
    for (int i = 0; i < n; ++i) Lp[i + 1] = Lp[i] + Lnz[i];
    Li.assign(Lp[n], 0);
    Lx.assign(Lp[n], 0.0);
    D.assign(n, 0.0);
    Dinv.assign(n, 0.0);
    mark.assign(n, 0);
    ylist.assign(n, 0);
    ebuf.assign(n, 0);
    nextcol.assign(n, 0);
    yv.assign(n, 0.0);
    return true;
  }

  bool numeric(const Csc& K) {
    for (int i = 0; i < n; ++i) {
      mark[i] = 0;
      yv[i] = 0.0;
      D[i] = 0.0;
      nextcol[i] = Lp[i];
    }
    // column 0 holds only its diagonal
    D[0] = K.x[0];
    if (D[0] == 0.0) return false;
    Dinv[0] = 1.0 / D[0];
    for (int k = 1; k < n; ++k) {
      int ny = 0;
      for (int q = K.p[k]; q < K.p[k + 1]; ++q) {
        int b = K.i[q];
        if (b == k) {
          D[k] = K.x[q];
          continue;
        }
        yv[b] = K.x[q];
        int nx = b;
        if (!mark[nx]) {
          mark[nx] = 1;
          ebuf[0] = nx;
          int ne = 1;
          nx = etree[b];
          while (nx != -1 && nx < k) {
            if (mark[nx]) break;
            mark[nx] = 1;
            ebuf[ne++] = nx;
            nx = etree[nx];
          }
          while (ne) ylist[ny++] = ebuf[--ne];
        }
      }
      for (int t = ny - 1; t >= 0; --t) {
        int c = ylist[t];
        int end = nextcol[c];
        double yc = yv[c];
        for (int j = Lp[c]; j < end; ++j) yv[Li[j]] -= Lx[j] * yc;
        Li[end] = k;
        Lx[end] = yc * Dinv[c];
        D[k] -= yc * Lx[end];
        nextcol[c]++;
        yv[c] = 0.0;
        mark[c] = 0;
      }
      if (D[k] == 0.0) return false;
      Dinv[k] = 1.0 / D[k];
    }
    return true;
  }

  void solve(double* x) const {
    for (int i = 0; i < n; ++i) {
      double v = x[i];
      for (int j = Lp[i]; j < Lp[i + 1]; ++j) x[Li[j]] -= Lx[j] * v;
    }
    for (int i = 0; i < n; ++i) x[i] *= Dinv[i];
    for (int i = n - 1; i >= 0; --i) {
      double v = x[i];
      for (int j = Lp[i]; j < Lp[i + 1]; ++j) v -= Lx[j] * x[Li[j]];
      x[i] = v;
    }
  }
};

// Plain minimum-degree ordering of a symmetric pattern (stand-in for SuiteSparse AMD, which
// OSQP's qdldl_interface.c calls; the ordering only changes round-off).  Returns perm with
// perm[new] = old.
inline std::vector<int> min_degree_order(int N, const std::vector<std::pair<int, int>>& edges) {
  std::vector<std::set<int>> adj(N);
  for (auto& e : edges)
    if (e.first != e.second) {
      adj[e.first].insert(e.second);
      adj[e.second].insert(e.first);
    }
  std::vector<char> done(N, 0);
  std::vector<int> perm;
  perm.reserve(N);
  for (int step = 0; step < N; ++step) {
    int best = -1;
    size_t bd = (size_t)-1;
    for (int v = 0; v < N; ++v)
      if (!done[v] && adj[v].size() < bd) {
        bd = adj[v].size();
        best = v;
      }
    done[best] = 1;
    perm.push_back(best);
    std::vector<int> nb(adj[best].begin(), adj[best].end());
    for (int a : nb) adj[a].erase(best);
    for (size_t s = 0; s < nb.size(); ++s)
      for (size_t t = s + 1; t < nb.size(); ++t) {
        adj[nb[s]].insert(nb[t]);
        adj[nb[t]].insert(nb[s]);
      }
    adj[best].clear();
  }
  return perm;
}

// ----------------------------------------------------------------------------------------
// OSQP v0.6.0 restated.
// ----------------------------------------------------------------------------------------
constexpr double RHO_MIN = 1e-06, RHO_MAX = 1e06, RHO_EQ_OVER_RHO_INEQ = 1e03, RHO_TOL = 1e-04;
constexpr double MIN_SCALING = 1e-04, MAX_SCALING = 1e+04;
constexpr double OSQP_INFTY = 1e30;
enum Status {
  DUAL_INFEASIBLE_INACCURATE = 4,
  PRIMAL_INFEASIBLE_INACCURATE = 3,
  SOLVED_INACCURATE = 2,
  SOLVED = 1,
  MAX_ITER_REACHED = -2,
  PRIMAL_INFEASIBLE = -3,
  DUAL_INFEASIBLE = -4,
  NON_CVX = -7,
  UNSOLVED = -10
};

struct Settings {  // osqp_set_default_settings (constants.h)
  double rho = 0.1, sigma = 1e-06;
  int scaling = 10;
  int adaptive_rho = 1, adaptive_rho_interval = 0;
  double adaptive_rho_tolerance = 5.0, adaptive_rho_fraction = 0.4;
  int max_iter = 4000;
  double eps_abs = 1e-3, eps_rel = 1e-3, eps_prim_inf = 1e-4, eps_dual_inf = 1e-4, alpha = 1.6;
  int scaled_termination = 0, check_termination = 25, warm_start = 1;
};

struct Osqp {
  int n = 0, m = 0;
  Csc P, A;  // scaled in place, as OSQP does
  std::vector<double> q, l, u;
  Settings st;
  // scaling
  double c = 1.0, cinv = 1.0;
  std::vector<double> D, Dinv, E, Einv, D_temp, D_temp_A, E_temp;
  // rho
  std::vector<double> rho_vec, rho_inv_vec;
  std::vector<int> constr_type;
  // iterates
  std::vector<double> x, z, y, x_prev, z_prev, xz_tilde, delta_x, delta_y, Ax, Px, Aty, Adelta_x,
      Atdelta_y, Pdelta_x;
  // KKT
  Csc K;                                      // permuted upper triangle
  std::vector<int> perm, pinv;                // perm[new]=old
  std::vector<int> kP, kA, kRho, kSig;        // positions in K.x of each source entry
  std::vector<double> sol, bp;
  Ldl ldl;
  // info
  int iter = 0, status = UNSOLVED, rho_updates = 0, n_factor = 0;
  double pri_res = 0, dua_res = 0;
  std::vector<double> sol_x, sol_y;
  // optional trace of (iter, pri_res, dua_res, rho) at every residual evaluation
  std::vector<double> trace;
  bool keep_trace = false;

  // ---------------------------------------------------------------- scaling.c
  static void limit_scaling(double* d, int len) {
    for (int i = 0; i < len; ++i) {
      d[i] = d[i] < MIN_SCALING ? 1.0 : d[i];
      d[i] = d[i] > MAX_SCALING ? MAX_SCALING : d[i];
    }
  }
  void scale_data() {
    c = 1.0;
    std::fill(D.begin(), D.end(), 1.0);
    std::fill(Dinv.begin(), Dinv.end(), 1.0);
    std::fill(E.begin(), E.end(), 1.0);
    std::fill(Einv.begin(), Einv.end(), 1.0);
    for (int it = 0; it < st.scaling; ++it) {
      // compute_inf_norm_cols_KKT
      mat_inf_norm_cols_sym_triu(P, D_temp.data());
      mat_inf_norm_cols(A, D_temp_A.data());
      for (int j = 0; j < n; ++j) D_temp[j] = std::max(D_temp[j], D_temp_A[j]);
      mat_inf_norm_rows(A, E_temp.data());
      limit_scaling(D_temp.data(), n);
      limit_scaling(E_temp.data(), m);
      for (int j = 0; j < n; ++j) D_temp[j] = std::sqrt(D_temp[j]);
      for (int i = 0; i < m; ++i) E_temp[i] = std::sqrt(E_temp[i]);
      for (int j = 0; j < n; ++j) D_temp[j] = 1.0 / D_temp[j];
      for (int i = 0; i < m; ++i) E_temp[i] = 1.0 / E_temp[i];
      mat_premult_diag(P, D_temp.data());
      mat_postmult_diag(P, D_temp.data());
      mat_premult_diag(A, E_temp.data());
      mat_postmult_diag(A, D_temp.data());
      for (int j = 0; j < n; ++j) q[j] = D_temp[j] * q[j];
      for (int j = 0; j < n; ++j) D[j] = D[j] * D_temp[j];
      for (int i = 0; i < m; ++i) E[i] = E[i] * E_temp[i];
      // cost normalisation
      mat_inf_norm_cols_sym_triu(P, D_temp.data());
      double c_temp = 0.0;
      for (int j = 0; j < n; ++j) c_temp += D_temp[j];
      c_temp /= (double)n;
      double inf_norm_q = vec_norm_inf(q.data(), n);
      limit_scaling(&inf_norm_q, 1);
      c_temp = std::max(c_temp, inf_norm_q);
      limit_scaling(&c_temp, 1);
      c_temp = 1.0 / c_temp;
      for (auto& v : P.x) v *= c_temp;
      for (int j = 0; j < n; ++j) q[j] *= c_temp;
      c *= c_temp;
    }
    cinv = 1.0 / c;
    for (int j = 0; j < n; ++j) Dinv[j] = 1.0 / D[j];
    for (int i = 0; i < m; ++i) Einv[i] = 1.0 / E[i];
    for (int i = 0; i < m; ++i) l[i] = E[i] * l[i];
    for (int i = 0; i < m; ++i) u[i] = E[i] * u[i];
  }
  void unscale_data() {
    for (auto& v : P.x) v *= cinv;
    mat_premult_diag(P, Dinv.data());
    mat_postmult_diag(P, Dinv.data());
    for (int j = 0; j < n; ++j) q[j] *= cinv;
    for (int j = 0; j < n; ++j) q[j] = Dinv[j] * q[j];
    mat_premult_diag(A, Einv.data());
    mat_postmult_diag(A, Dinv.data());
    for (int i = 0; i < m; ++i) l[i] = Einv[i] * l[i];
    for (int i = 0; i < m; ++i) u[i] = Einv[i] * u[i];
  }

  // ---------------------------------------------------------------- auxil.c: rho handling
  void set_rho_vec() {
    st.rho = std::min(std::max(st.rho, RHO_MIN), RHO_MAX);
    for (int i = 0; i < m; ++i) {
      if ((l[i] < -OSQP_INFTY * MIN_SCALING) && (u[i] > OSQP_INFTY * MIN_SCALING)) {
        constr_type[i] = -1;
        rho_vec[i] = RHO_MIN;
      } else if (u[i] - l[i] < RHO_TOL) {
        constr_type[i] = 1;
        rho_vec[i] = RHO_EQ_OVER_RHO_INEQ * st.rho;
      } else {
        constr_type[i] = 0;
        rho_vec[i] = st.rho;
      }
      rho_inv_vec[i] = 1.0 / rho_vec[i];
    }
  }
  void update_rho_vec() {
    bool changed = false;
    for (int i = 0; i < m; ++i) {
      if ((l[i] < -OSQP_INFTY * MIN_SCALING) && (u[i] > OSQP_INFTY * MIN_SCALING)) {
        if (constr_type[i] != -1) {
          constr_type[i] = -1;
          rho_vec[i] = RHO_MIN;
          rho_inv_vec[i] = 1.0 / RHO_MIN;
          changed = true;
        }
      } else if (u[i] - l[i] < RHO_TOL) {
        if (constr_type[i] != 1) {
          constr_type[i] = 1;
          rho_vec[i] = RHO_EQ_OVER_RHO_INEQ * st.rho;
          rho_inv_vec[i] = 1.0 / rho_vec[i];
          changed = true;
        }
      } else {
        if (constr_type[i] != 0) {
          constr_type[i] = 0;
          rho_vec[i] = st.rho;
          rho_inv_vec[i] = 1.0 / st.rho;
          changed = true;
        }
      }
    }
    if (changed) refactor();
  }
  void osqp_update_rho(double rho_new) {
    if (rho_new <= 0) return;
    st.rho = std::min(std::max(rho_new, RHO_MIN), RHO_MAX);
    for (int i = 0; i < m; ++i) {
      if (constr_type[i] == 0) {
        rho_vec[i] = st.rho;
        rho_inv_vec[i] = 1.0 / st.rho;
      } else if (constr_type[i] == 1) {
        rho_vec[i] = RHO_EQ_OVER_RHO_INEQ * st.rho;
        rho_inv_vec[i] = 1.0 / rho_vec[i];
      }
    }
    refactor();
  }

  // ---------------------------------------------------------------- kkt.c / qdldl_interface.c
  void build_kkt_pattern() {
    const int N = n + m;
    // source triplets in original indexing (upper triangle)
    struct Ent {
      int i, j, kind, src;
    };  // kind 0=P 1=A 2=rho 3=sigma
    std::vector<Ent> ent;
    std::vector<std::pair<int, int>> edges;
    for (int j = 0; j < n; ++j)
      for (int k = P.p[j]; k < P.p[j + 1]; ++k) ent.push_back({P.i[k], j, 0, k});
    for (int j = 0; j < n; ++j) ent.push_back({j, j, 3, j});
    for (int j = 0; j < n; ++j)
      for (int k = A.p[j]; k < A.p[j + 1]; ++k) ent.push_back({j, n + A.i[k], 1, k});
    for (int i = 0; i < m; ++i) ent.push_back({n + i, n + i, 2, i});
    for (auto& e : ent) edges.push_back({e.i, e.j});
    perm = min_degree_order(N, edges);
    pinv.assign(N, 0);
    for (int k = 0; k < N; ++k) pinv[perm[k]] = k;
    // permuted upper-triangular pattern
    std::vector<int> ord(ent.size());
    std::vector<int> pi(ent.size()), pj(ent.size());
    for (size_t e = 0; e < ent.size(); ++e) {
      int a = pinv[ent[e].i], b = pinv[ent[e].j];
      pi[e] = std::min(a, b);
      pj[e] = std::max(a, b);
      ord[e] = (int)e;
    }
    std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) {
      if (pj[a] != pj[b]) return pj[a] < pj[b];
      return pi[a] < pi[b];
    });
    K = Csc();
    K.m = K.n = N;
    K.p.assign(N + 1, 0);
    kP.assign(P.nnz(), -1);
    kA.assign(A.nnz(), -1);
    kRho.assign(m, -1);
    kSig.assign(n, -1);
    int lr = -1, lc = -1;
    for (size_t t = 0; t < ord.size(); ++t) {
      int e = ord[t];
      if (!(pi[e] == lr && pj[e] == lc)) {
        K.i.push_back(pi[e]);
        K.x.push_back(0.0);
        K.p[pj[e] + 1]++;
        lr = pi[e];
        lc = pj[e];
      }
      int pos = (int)K.x.size() - 1;
      switch (ent[e].kind) {
        case 0: kP[ent[e].src] = pos; break;
        case 1: kA[ent[e].src] = pos; break;
        case 2: kRho[ent[e].src] = pos; break;
        default: kSig[ent[e].src] = pos; break;
      }
    }
    for (int j = 0; j < N; ++j) K.p[j + 1] += K.p[j];
    ldl.symbolic(K);
    sol.assign(N, 0.0);
    bp.assign(N, 0.0);
  }
  void refactor() {
    std::fill(K.x.begin(), K.x.end(), 0.0);
    for (size_t k = 0; k < kP.size(); ++k) K.x[kP[k]] += P.x[k];
    for (int j = 0; j < n; ++j) K.x[kSig[j]] += st.sigma;
    for (size_t k = 0; k < kA.size(); ++k) K.x[kA[k]] += A.x[k];
    for (int i = 0; i < m; ++i) K.x[kRho[i]] += -rho_inv_vec[i];
    ldl.numeric(K);
    n_factor++;
  }
  // solve_linsys_qdldl: b = [rhs_x ; rhs_z] in, [x_tilde ; z_tilde] out
  void kkt_solve(double* b) {
    const int N = n + m;
    for (int k = 0; k < N; ++k) bp[k] = b[perm[k]];
    ldl.solve(bp.data());
    for (int k = 0; k < N; ++k) sol[perm[k]] = bp[k];
    for (int j = 0; j < n; ++j) b[j] = sol[j];
    for (int j = 0; j < m; ++j) b[j + n] += rho_inv_vec[j] * sol[j + n];
  }

  // ---------------------------------------------------------------- osqp.c: setup / updates
  void setup(const Csc& P_in, const Csc& A_in, const double* q_in, const double* l_in,
             const double* u_in, const Settings& s) {
    n = A_in.n;
    m = A_in.m;
    P = P_in;
    A = A_in;
    q.assign(q_in, q_in + n);
    l.assign(l_in, l_in + m);
    u.assign(u_in, u_in + m);
    st = s;
    D.assign(n, 1.0);
    Dinv.assign(n, 1.0);
    E.assign(m, 1.0);
    Einv.assign(m, 1.0);
    D_temp.assign(n, 0.0);
    D_temp_A.assign(n, 0.0);
    E_temp.assign(m, 0.0);
    rho_vec.assign(m, 0.0);
    rho_inv_vec.assign(m, 0.0);
    constr_type.assign(m, 0);
    x.assign(n, 0.0);
    z.assign(m, 0.0);
    y.assign(m, 0.0);
    x_prev.assign(n, 0.0);
    z_prev.assign(m, 0.0);
    xz_tilde.assign(n + m, 0.0);
    delta_x.assign(n, 0.0);
    delta_y.assign(m, 0.0);
    Ax.assign(m, 0.0);
    Px.assign(n, 0.0);
    Aty.assign(n, 0.0);
    Adelta_x.assign(m, 0.0);
    Atdelta_y.assign(n, 0.0);
    Pdelta_x.assign(n, 0.0);
    sol_x.assign(n, 0.0);
    sol_y.assign(m, 0.0);
    if (st.scaling) scale_data();
    set_rho_vec();
    build_kkt_pattern();
    refactor();
    if (st.adaptive_rho && !st.adaptive_rho_interval)
      st.adaptive_rho_interval = st.check_termination ? 4 * st.check_termination : 100;
    status = UNSOLVED;
    rho_updates = 0;
  }
  void reset_info() {
    status = UNSOLVED;
    rho_updates = 0;
  }
  void update_A(const double* Ax_new) {
    if (st.scaling) unscale_data();
    for (int k = 0; k < A.nnz(); ++k) A.x[k] = Ax_new[k];
    if (st.scaling) scale_data();
    refactor();
    reset_info();
  }
  void update_P(const double* Px_new) {
    if (st.scaling) unscale_data();
    for (int k = 0; k < P.nnz(); ++k) P.x[k] = Px_new[k];
    if (st.scaling) scale_data();
    refactor();
    reset_info();
  }
  void update_lin_cost(const double* q_new) {
    for (int j = 0; j < n; ++j) q[j] = q_new[j];
    if (st.scaling) {
      for (int j = 0; j < n; ++j) q[j] = D[j] * q[j];
      for (int j = 0; j < n; ++j) q[j] *= c;
    }
    reset_info();
  }
  int update_bounds(const double* l_new, const double* u_new) {
    for (int i = 0; i < m; ++i)
      if (l_new[i] > u_new[i]) return 1;
    for (int i = 0; i < m; ++i) l[i] = l_new[i];
    for (int i = 0; i < m; ++i) u[i] = u_new[i];
    if (st.scaling) {
      for (int i = 0; i < m; ++i) l[i] = E[i] * l[i];
      for (int i = 0; i < m; ++i) u[i] = E[i] * u[i];
    }
    reset_info();
    update_rho_vec();
    return 0;
  }
  int update_upper_bound(const double* u_new) {
    for (int i = 0; i < m; ++i) u[i] = u_new[i];
    if (st.scaling)
      for (int i = 0; i < m; ++i) u[i] = E[i] * u[i];
    for (int i = 0; i < m; ++i)
      if (u[i] < l[i]) return 1;
    reset_info();
    update_rho_vec();
    return 0;
  }
  int update_lower_bound(const double* l_new) {
    for (int i = 0; i < m; ++i) l[i] = l_new[i];
    if (st.scaling)
      for (int i = 0; i < m; ++i) l[i] = E[i] * l[i];
    for (int i = 0; i < m; ++i)
      if (l[i] > u[i]) return 1;
    reset_info();
    update_rho_vec();
    return 0;
  }

  // ---------------------------------------------------------------- auxil.c: residuals
  double compute_pri_res() {
    mat_vec(A, x.data(), Ax.data(), 0);
    for (int i = 0; i < m; ++i) z_prev[i] = Ax[i] - z[i];  // z_prev is OSQP's work vector here
    if (st.scaling && !st.scaled_termination)
      return vec_scaled_norm_inf(Einv.data(), z_prev.data(), m);
    return vec_norm_inf(z_prev.data(), m);
  }
  double compute_dua_res() {
    for (int j = 0; j < n; ++j) x_prev[j] = q[j];  // x_prev is OSQP's work vector here
    mat_vec(P, x.data(), Px.data(), 0);
    mat_tpose_vec(P, x.data(), Px.data(), 1, 1);
    for (int j = 0; j < n; ++j) x_prev[j] = x_prev[j] + Px[j];
    if (m > 0) {
      mat_tpose_vec(A, y.data(), Aty.data(), 0, 0);
      for (int j = 0; j < n; ++j) x_prev[j] = x_prev[j] + Aty[j];
    }
    if (st.scaling && !st.scaled_termination)
      return cinv * vec_scaled_norm_inf(Dinv.data(), x_prev.data(), n);
    return vec_norm_inf(x_prev.data(), n);
  }
  void update_info(int it) {
    iter = it;
    pri_res = (m == 0) ? 0.0 : compute_pri_res();
    dua_res = compute_dua_res();
    if (keep_trace) {
      trace.push_back((double)it);
      trace.push_back(pri_res);
      trace.push_back(dua_res);
      trace.push_back(st.rho);
    }
  }
  double compute_pri_tol(double eps_abs, double eps_rel) {
    double mx;
    if (st.scaling && !st.scaled_termination) {
      mx = vec_scaled_norm_inf(Einv.data(), z.data(), m);
      mx = std::max(mx, vec_scaled_norm_inf(Einv.data(), Ax.data(), m));
    } else {
      mx = std::max(vec_norm_inf(z.data(), m), vec_norm_inf(Ax.data(), m));
    }
    return eps_abs + eps_rel * mx;
  }
  double compute_dua_tol(double eps_abs, double eps_rel) {
    double mx;
    if (st.scaling && !st.scaled_termination) {
      mx = vec_scaled_norm_inf(Dinv.data(), q.data(), n);
      mx = std::max(mx, vec_scaled_norm_inf(Dinv.data(), Aty.data(), n));
      mx = std::max(mx, vec_scaled_norm_inf(Dinv.data(), Px.data(), n));
      mx *= cinv;
    } else {
      mx = vec_norm_inf(q.data(), n);
      mx = std::max(mx, vec_norm_inf(Aty.data(), n));
      mx = std::max(mx, vec_norm_inf(Px.data(), n));
    }
    return eps_abs + eps_rel * mx;
  }
  bool is_primal_infeasible(double eps_prim_inf) {
    for (int i = 0; i < m; ++i) {
      if (u[i] > OSQP_INFTY * MIN_SCALING) {
        if (l[i] < -OSQP_INFTY * MIN_SCALING)
          delta_y[i] = 0.0;
        else
          delta_y[i] = std::min(delta_y[i], 0.0);
      } else if (l[i] < -OSQP_INFTY * MIN_SCALING) {
        delta_y[i] = std::max(delta_y[i], 0.0);
      }
    }
    double norm_delta_y;
    if (st.scaling && !st.scaled_termination) {
      for (int i = 0; i < m; ++i) Adelta_x[i] = E[i] * delta_y[i];
      norm_delta_y = vec_norm_inf(Adelta_x.data(), m);
    } else {
      norm_delta_y = vec_norm_inf(delta_y.data(), m);
    }
    if (norm_delta_y > eps_prim_inf) {
      double ineq_lhs = 0.0;
      // NB: with l = -inf (the MPC's cone rows, src/MPC.cpp:293-300) this sum is NaN
      // (-inf * 0), every comparison below is false and the test can never fire -- exactly
      // what the C code does with IEEE doubles.
      for (int i = 0; i < m; ++i)
        ineq_lhs += u[i] * std::max(delta_y[i], 0.0) + l[i] * std::min(delta_y[i], 0.0);
      if (ineq_lhs < -eps_prim_inf * norm_delta_y) {
        mat_tpose_vec(A, delta_y.data(), Atdelta_y.data(), 0, 0);
        if (st.scaling && !st.scaled_termination)
          for (int j = 0; j < n; ++j) Atdelta_y[j] = Dinv[j] * Atdelta_y[j];
        return vec_norm_inf(Atdelta_y.data(), n) < eps_prim_inf * norm_delta_y;
      }
    }
    return false;
  }
  bool is_dual_infeasible(double eps_dual_inf) {
    double norm_delta_x, cost_scaling;
    if (st.scaling && !st.scaled_termination) {
      norm_delta_x = vec_scaled_norm_inf(D.data(), delta_x.data(), n);
      cost_scaling = c;
    } else {
      norm_delta_x = vec_norm_inf(delta_x.data(), n);
      cost_scaling = 1.0;
    }
    if (norm_delta_x > eps_dual_inf) {
      double qdx = 0.0;
      for (int j = 0; j < n; ++j) qdx += q[j] * delta_x[j];
      if (qdx < -cost_scaling * eps_dual_inf * norm_delta_x) {
        mat_vec(P, delta_x.data(), Pdelta_x.data(), 0);
        mat_tpose_vec(P, delta_x.data(), Pdelta_x.data(), 1, 1);
        if (st.scaling && !st.scaled_termination)
          for (int j = 0; j < n; ++j) Pdelta_x[j] = Dinv[j] * Pdelta_x[j];
        if (vec_norm_inf(Pdelta_x.data(), n) < cost_scaling * eps_dual_inf * norm_delta_x) {
          mat_vec(A, delta_x.data(), Adelta_x.data(), 0);
          if (st.scaling && !st.scaled_termination)
            for (int i = 0; i < m; ++i) Adelta_x[i] = Einv[i] * Adelta_x[i];
          for (int i = 0; i < m; ++i) {
            if (((u[i] < OSQP_INFTY * MIN_SCALING) &&
                 (Adelta_x[i] > eps_dual_inf * norm_delta_x)) ||
                ((l[i] > -OSQP_INFTY * MIN_SCALING) &&
                 (Adelta_x[i] < -eps_dual_inf * norm_delta_x)))
              return false;
          }
          return true;
        }
      }
    }
    return false;
  }
  bool check_termination(int approximate) {
    double eps_abs = st.eps_abs, eps_rel = st.eps_rel, eps_prim_inf = st.eps_prim_inf,
           eps_dual_inf = st.eps_dual_inf;
    bool prim_res_check = false, dual_res_check = false, prim_inf_check = false,
         dual_inf_check = false;
    if ((pri_res > OSQP_INFTY) || (dua_res > OSQP_INFTY)) {
      status = NON_CVX;
      return true;
    }
    if (approximate) {
      eps_abs *= 10;
      eps_rel *= 10;
      eps_prim_inf *= 10;
      eps_dual_inf *= 10;
    }
    if (m == 0) {
      prim_res_check = true;
    } else {
      double eps_prim = compute_pri_tol(eps_abs, eps_rel);
      if (pri_res < eps_prim)
        prim_res_check = true;
      else
        prim_inf_check = is_primal_infeasible(eps_prim_inf);
    }
    double eps_dual = compute_dua_tol(eps_abs, eps_rel);
    if (dua_res < eps_dual)
      dual_res_check = true;
    else
      dual_inf_check = is_dual_infeasible(eps_dual_inf);
    if (prim_res_check && dual_res_check) {
      status = approximate ? SOLVED_INACCURATE : SOLVED;
      return true;
    } else if (prim_inf_check) {
      status = approximate ? PRIMAL_INFEASIBLE_INACCURATE : PRIMAL_INFEASIBLE;
      if (st.scaling && !st.scaled_termination)
        for (int i = 0; i < m; ++i) delta_y[i] = E[i] * delta_y[i];
      return true;
    } else if (dual_inf_check) {
      status = approximate ? DUAL_INFEASIBLE_INACCURATE : DUAL_INFEASIBLE;
      if (st.scaling && !st.scaled_termination)
        for (int j = 0; j < n; ++j) delta_x[j] = D[j] * delta_x[j];
      return true;
    }
    return false;
  }
  double compute_rho_estimate() {
    double pr = vec_norm_inf(z_prev.data(), m);  // residual vectors left in the work buffers
    double dr = vec_norm_inf(x_prev.data(), n);
    double pn = std::max(vec_norm_inf(z.data(), m), vec_norm_inf(Ax.data(), m));
    pr /= (pn + 1e-10);
    double dn = vec_norm_inf(q.data(), n);
    dn = std::max(dn, vec_norm_inf(Aty.data(), n));
    dn = std::max(dn, vec_norm_inf(Px.data(), n));
    dr /= (dn + 1e-10);
    double est = st.rho * std::sqrt(pr / (dr + 1e-10));
    return std::min(std::max(est, RHO_MIN), RHO_MAX);
  }
  void adapt_rho() {
    double rho_new = compute_rho_estimate();
    if ((rho_new > st.rho * st.adaptive_rho_tolerance) ||
        (rho_new < st.rho / st.adaptive_rho_tolerance)) {
      osqp_update_rho(rho_new);
      rho_updates++;
    }
  }

  // ---------------------------------------------------------------- osqp.c: osqp_solve
  int solve() {
    bool can_check = false;
    int it;
    if (keep_trace) trace.clear();
    if (!st.warm_start) {
      std::fill(x.begin(), x.end(), 0.0);
      std::fill(z.begin(), z.end(), 0.0);
      std::fill(y.begin(), y.end(), 0.0);
    }
    for (it = 1; it <= st.max_iter; ++it) {
      x.swap(x_prev);
      z.swap(z_prev);
      // update_xz_tilde
      for (int j = 0; j < n; ++j) xz_tilde[j] = st.sigma * x_prev[j] - q[j];
      for (int i = 0; i < m; ++i) xz_tilde[i + n] = z_prev[i] - rho_inv_vec[i] * y[i];
      kkt_solve(xz_tilde.data());
      // update_x
      for (int j = 0; j < n; ++j) x[j] = st.alpha * xz_tilde[j] + (1.0 - st.alpha) * x_prev[j];
      for (int j = 0; j < n; ++j) delta_x[j] = x[j] - x_prev[j];
      // update_z
      for (int i = 0; i < m; ++i)
        z[i] = st.alpha * xz_tilde[i + n] + (1.0 - st.alpha) * z_prev[i] + rho_inv_vec[i] * y[i];
      for (int i = 0; i < m; ++i) z[i] = std::min(std::max(z[i], l[i]), u[i]);
      // update_y
      for (int i = 0; i < m; ++i) {
        delta_y[i] =
            rho_vec[i] * (st.alpha * xz_tilde[i + n] + (1.0 - st.alpha) * z_prev[i] - z[i]);
        y[i] += delta_y[i];
      }
      can_check = st.check_termination && (it % st.check_termination == 0);
      if (can_check) {
        update_info(it);
        if (check_termination(0)) break;
      }
      if (st.adaptive_rho && st.adaptive_rho_interval && (it % st.adaptive_rho_interval == 0)) {
        if (!can_check) update_info(it);
        adapt_rho();
      }
    }
    if (!can_check) {
      update_info(it - 1);
      check_termination(0);
    }
    if (status == UNSOLVED) {
      if (!check_termination(1)) status = MAX_ITER_REACHED;
    }
    // store_solution
    bool has_sol = (status != PRIMAL_INFEASIBLE) && (status != PRIMAL_INFEASIBLE_INACCURATE) &&
                   (status != DUAL_INFEASIBLE) && (status != DUAL_INFEASIBLE_INACCURATE) &&
                   (status != NON_CVX);
    if (has_sol) {
      for (int j = 0; j < n; ++j) sol_x[j] = x[j];
      for (int i = 0; i < m; ++i) sol_y[i] = y[i];
      if (st.scaling) {
        for (int j = 0; j < n; ++j) sol_x[j] = D[j] * sol_x[j];
        for (int i = 0; i < m; ++i) sol_y[i] = E[i] * sol_y[i];
        for (int i = 0; i < m; ++i) sol_y[i] *= cinv;
      }
    } else {
      const double nan = std::numeric_limits<double>::quiet_NaN();
      std::fill(sol_x.begin(), sol_x.end(), nan);
      std::fill(sol_y.begin(), sol_y.end(), nan);
      std::fill(x.begin(), x.end(), 0.0);
      std::fill(z.begin(), z.end(), 0.0);
      std::fill(y.begin(), y.end(), 0.0);
    }
    return 0;
  }
};

}  // namespace orc
