// tests/emu/emu_planner.cpp -- TEST INFRASTRUCTURE.  The product's planner core (csrc/planner_core.cuh) compiled by a
// plain C++ compiler (-DQRW_EMULATE): the four lanes of a robot run one after the other, every lane's lane_compute before
// any lane's lane_store (what __syncwarp orders on the GPU).  Never loaded by the product; not a CPU fallback.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../quadruped-reactive-walking_b200/csrc/planner_core.cuh"

using namespace qrw::plan;

namespace {
struct EmuPlanner {
  PlanConst C;
  PlanMem M;
  std::vector<double> st, xref, fsteps;
  std::vector<unsigned long long> bits;
  std::vector<int> flags;
};
int g_lane_order = 0;
}  // namespace

extern "C" {

void* emu_planner_create(int B, double dt_mpc, double dt_wbc, double T_gait, double T_mpc, int N_gait, int k_mpc, double h_ref,
                         const double* shoulders, const double* ftg_target, const double* ftg_pos, double max_height,
                         double lock_time) {
  auto* P = new EmuPlanner();
  PlanConst& C = P->C;
  const int n = (int)std::lround(T_mpc / dt_mpc);
  C.dt_mpc = dt_mpc, C.dt_wbc = dt_wbc, C.T_gait = T_gait, C.T_mpc = T_mpc, C.h_ref = h_ref, C.max_height = max_height,
  C.lock_time = lock_time, C.lin_step = n > 1 ? (T_mpc - dt_mpc) / (n - 1) : 0.0;
  for (int i = 0; i < 12; ++i) C.shoulders[i] = shoulders[i];
  C.N_gait = N_gait, C.n_steps = n, C.k_mpc = k_mpc, C.pad = 0;
  P->st.assign((size_t)PLAN_STATE_DOUBLES * B, 0.0);
  P->xref.assign((size_t)B * 12 * (n + 1), 0.0);
  P->fsteps.assign((size_t)B * 12 * N_gait, 0.0);
  P->bits.assign((size_t)6 * B, 0ull);
  P->flags.assign(B, 2);
  P->M = PlanMem{B, P->st.data(), P->bits.data(), P->flags.data(), P->xref.data(), P->fsteps.data()};
  GaitState g;
  g.past.clear(), g.cur.clear(), g.des.clear();
  g.newPhase = 0, g.is_static = 1, g.remaining = 0.0;
  g.create(C, 3);
  g.create_gait_f(C);
  const GaitBits* m[3] = {&g.past, &g.cur, &g.des};
  for (int f = 0; f < 3; ++f)
    for (int b = 0; b < B; ++b) P->bits[((size_t)f * B + b) * 2] = m[f]->lo, P->bits[((size_t)f * B + b) * 2 + 1] = m[f]->hi;
  for (int b = 0; b < B; ++b) {
    for (int i = 0; i < 12; ++i) {
      P->M.f(F_CURFS, 12, b)[i] = P->M.f(F_TARGET, 12, b)[i] = P->M.f(F_OTARGET, 12, b)[i] = shoulders[i];
      P->M.f(F_FTGTGT, 12, b)[i] = ftg_target[i];
      P->M.f(F_FPOS, 12, b)[i] = ftg_pos[i];
    }
    for (int j = 0; j < 4; ++j) P->M.f(F_CONTACTS, 4, b)[j] = g.cur.at(0, j) ? 1.0 : 0.0;
  }
  return P;
}
void emu_planner_destroy(void* h) { delete static_cast<EmuPlanner*>(h); }

void emu_planner_run(void* h, int k, int stages, const double* q7, const double* v6, const double* vref6, const int* code) {
  auto* P = static_cast<EmuPlanner*>(h);
  const int N = P->C.N_gait;
  for (int b = 0; b < P->M.B; ++b) {
    LaneCtx X[4];
    for (int jj = 0; jj < 4; ++jj) {
      const int j = g_lane_order ? 3 - jj : jj;
      lane_compute(P->C, P->M, b, j, k, stages, q7 + 7 * b, v6 + 6 * b, vref6 + 6 * b, code ? code[b] : 0,
                   P->fsteps.data() + (size_t)b * N * 12 + 3 * N * j, X[j]);
    }
    for (int jj = 0; jj < 4; ++jj) {
      const int j = g_lane_order == 1 ? 3 - jj : jj;
      lane_store(P->C, P->M, b, j, X[j]);
    }
  }
}
// 0: lanes 0..3 in both phases, 1: 3..0 in both, 2: compute 3..0, store 0..3 (every order is a legal GPU schedule)
void emu_planner_set_lane_order(int order) { g_lane_order = order; }

// which: 0 xref, 1 fsteps, 2..: a [B][len] field of the state (field offset = which - 2)
void emu_planner_get(void* h, int which, int len, double* out) {
  auto* P = static_cast<EmuPlanner*>(h);
  if (which == 0) std::memcpy(out, P->xref.data(), P->xref.size() * sizeof(double));
  else if (which == 1) std::memcpy(out, P->fsteps.data(), P->fsteps.size() * sizeof(double));
  else std::memcpy(out, P->st.data() + (size_t)P->M.B * (which - 2), (size_t)P->M.B * len * sizeof(double));
}
// m: 0 past, 1 current, 2 desired -> [B][N_gait*4] column-major doubles
void emu_planner_gait(void* h, int m, double* out) {
  auto* P = static_cast<EmuPlanner*>(h);
  const int N = P->C.N_gait, B = P->M.B;
  for (int b = 0; b < B; ++b) {
    GaitBits g{P->bits[((size_t)m * B + b) * 2], P->bits[((size_t)m * B + b) * 2 + 1]};
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < 4; ++j) out[(size_t)b * N * 4 + i + N * j] = g.at(i, j) ? 1.0 : 0.0;
  }
}
void emu_planner_flags(void* h, int* out) {  // [B][2] newPhase, is_static
  auto* P = static_cast<EmuPlanner*>(h);
  for (int b = 0; b < P->M.B; ++b) out[2 * b] = P->flags[b] & 1, out[2 * b + 1] = (P->flags[b] >> 1) & 1;
}
int emu_planner_field(const char* name) {
  struct { const char* n; int f; } t[] = {{"cur_fs", F_CURFS}, {"target", F_TARGET}, {"o_target", F_OTARGET}, {"ftg_target", F_FTGTGT},
      {"fpos", F_FPOS}, {"fvel", F_FVEL}, {"facc", F_FACC}, {"p_cmd", F_PCMD}, {"v_cmd", F_VCMD}, {"a_cmd", F_ACMD},
      {"t0s", F_T0S}, {"t_swing", F_TSW}, {"contacts", F_CONTACTS}};
  for (auto& e : t)
    if (!std::strcmp(e.n, name)) return e.f + 2;
  return -1;
}
}  // extern "C"
