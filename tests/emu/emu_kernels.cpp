// tests/emu/emu_kernels.cpp -- TEST INFRASTRUCTURE.  Compiles the product's kernel cores
// (csrc/*_core.cuh) with a plain C++ compiler in warp-emulation mode (-DQRW_EMULATE, see
// csrc/qrw_spmd.cuh) so their logic can be checked against the CPU oracle on a machine that has
// no GPU.  It is never loaded by the product and is not a CPU fallback.
#include <cstring>
#include <thread>
#include <vector>

#include "cta_emu.hpp"
#include "../../quadruped-reactive-walking_b200/csrc/mpc_cta_core.cuh"
#include "../../quadruped-reactive-walking_b200/csrc/mpc_v2_core.cuh"
#include "../../quadruped-reactive-walking_b200/csrc/mpc_v3_core.cuh"
#include "../../quadruped-reactive-walking_b200/csrc/mpc_v4_core.cuh"

using namespace qrw;

extern "C" {

int emu_mpc_state_doubles() { return MPC_STATE_DOUBLES; }
// 0 / 1 / 2: order in which the emulated threads of a CTA are run between collectives (cta_emu.hpp)
void emu_set_schedule(int order) { qrw::cta::g_sweep_order = order; }
// warp-SPMD cores (mpc_core.cuh, wbc_core.cuh): lanes of a phase ascending (0) or descending (1)
void emu_set_lanes_descending(int on) { qrw::g_emu_lanes_descending = on; }
int emu_mpc_state_ints() { return MPC_STATE_INTS; }

// Same contract as the CUDA kernel: instance-major inputs, persistent state, outputs.
void emu_mpc_run(int B, double dt, int N_gait, int first, int max_iter, const double* xref,
                 const double* fsteps, double* st_d, int* st_i, double* out, int* info, double* rho,
                 double* dbg, int nthreads) {
  MpcConst C = make_mpc_const(dt, N_gait);
  if (max_iter > 0) C.max_iter = max_iter;
  auto work = [&](int t) {
    std::vector<double> slab(MPC_SLAB_DOUBLES, 0.0);
    for (int b = t; b < B; b += nthreads)
      mpc_solve_instance(C, slab.data(), xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * N_gait,
                         st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, first,
                         out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr);
  };
  if (nthreads <= 1) {
    nthreads = 1;
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
    for (auto& t : th) t.join();
  }
}
}

#include "../../quadruped-reactive-walking_b200/csrc/wbc_core.cuh"
extern "C" {
int emu_wbc_state_doubles() { return WBC_STATE_DOUBLES; }
void emu_wbc_run(int B, int first, const double* M, const double* Jc, const double* f_cmd, const double* RNEA,
                 double* st, double* f_res, double* ddq_res, double* H, int* info) {
  WbcConst C = make_wbc_const();
  std::vector<double> slab(WBC_SLAB_DOUBLES, 0.0);
  for (int b = 0; b < B; ++b)
    wbc_solve_instance(C, slab.data(), M + (size_t)b * 324, Jc + (size_t)b * 216, f_cmd + 12 * b, RNEA + 6 * b,
                       st + (size_t)b * WBC_STATE_DOUBLES, first, f_res + 12 * b, ddq_res + 6 * b,
                       H ? H + 144 * b : nullptr, info + 4 * b);
}
void emu_invkin_run(int B, const double* contacts, const double* goals, const double* vgoals, const double* agoals,
                    const double* posf, const double* vf, const double* wf, const double* af, const double* Jf,
                    double* ddq, double* dq_cmd, double* q_step) {
  for (int b = 0; b < B; ++b)
    invkin_instance(contacts + 4 * b, goals + 12 * b, vgoals + 12 * b, agoals + 12 * b, posf + 12 * b, vf + 12 * b,
                    wf + 12 * b, af + 12 * b, Jf + 144 * b, ddq + 12 * b, dq_cmd + 12 * b, q_step + 12 * b);
}
}

static std::vector<double> g_last_slab;  // debugging aid: shared slab of the last emulated CTA
extern "C" int emu_last_slab(double* out, int cap) {
  int n = (int)g_last_slab.size();
  for (int i = 0; i < n && i < cap; ++i) out[i] = g_last_slab[i];
  return n;
}
#ifdef QRW_DEBUG_SOLVE
extern "C" void emu_dbgK(double* out) { memcpy(out, cta::g_dbgK, sizeof(cta::g_dbgK)); }
#endif
// CTA-per-instance MPC kernel core (csrc/mpc_cta_core.cuh): every thread of the 256-thread CTA is a fiber
// (cta_emu.hpp); same contract as emu_mpc_run.
extern "C" void emu_mpc_cta_run(int B, double dt, int N_gait, int first, int max_iter, const double* xref,
                                const double* fsteps, double* st_d, int* st_i, double* out, int* info, double* rho,
                                double* dbg, int nthreads) {
  MpcConst C = make_mpc_const(dt, N_gait);
  if (max_iter > 0) C.max_iter = max_iter;
  auto work = [&](int t) {
    std::vector<double> slab(cta::CTA_SLAB_DOUBLES, 0.0);
    struct Keep {
      std::vector<double>& s;
      ~Keep() { g_last_slab = s; }
    } keep{slab};
    for (int b = t; b < B; b += nthreads) {
      cta::run_cta(cta::NT, [&]() {
        cta::mpc_cta_solve(C, slab.data(), xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * N_gait,
                           st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, first,
                           out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr);
      });
    }
  };
  if (nthreads <= 1) {
    nthreads = 1;
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
    for (auto& t : th) t.join();
  }
}

// Force-first CTA kernel core (csrc/mpc_v2_core.cuh), every thread of the 256-thread CTA a fiber; same contract.
extern "C" void emu_mpc_v2_run(int B, double dt, int N_gait, int first, int max_iter, const double* xref,
                               const double* fsteps, double* st_d, int* st_i, double* out, int* info, double* rho,
                               double* dbg, int nthreads) {
  MpcConst C = make_mpc_const(dt, N_gait);
  if (max_iter > 0) C.max_iter = max_iter;
  auto work = [&](int t) {
    std::vector<double> slab(v2::V2_SLAB_DOUBLES, 0.0);
    for (int b = t; b < B; b += nthreads) {
      cta::run_cta(v2::NT, [&]() {
        v2::mpc_v2_solve(C, slab.data(), xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * N_gait,
                         st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, first,
                         out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr, b % 8, 0u);
      });
    }
    if (nthreads <= 1) g_last_slab = slab;
  };
  if (nthreads <= 1) {
    nthreads = 1;
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
    for (auto& t : th) t.join();
  }
}

// Super-stage chain CTA kernel core (csrc/mpc_v4_core.cuh), every thread of the 256-thread CTA a fiber; same contract.
extern "C" void emu_mpc_v4_run(int B, double dt, int N_gait, int first, int max_iter, const double* xref,
                               const double* fsteps, double* st_d, int* st_i, double* out, int* info, double* rho,
                               double* dbg, int nthreads) {
  MpcConst C = make_mpc_const(dt, N_gait);
  if (max_iter > 0) C.max_iter = max_iter;
  auto work = [&](int t) {
    std::vector<double> slab(v4::V4_SLAB_DOUBLES, 0.0);
    for (int b = t; b < B; b += nthreads) {
      cta::run_cta(v4::NT, [&]() {
        v4::mpc_v4_solve(C, slab.data(), xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * N_gait,
                         st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, first,
                         out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr, b % 8, 0u);
      });
    }
    if (nthreads <= 1) g_last_slab = slab;
  };
  if (nthreads <= 1) {
    nthreads = 1;
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
    for (auto& t : th) t.join();
  }
}

// Nested-dissection CTA kernel core (csrc/mpc_v3_core.cuh, A/B variant), every thread of the 256-thread CTA a fiber; same contract.
extern "C" void emu_mpc_v3_run(int B, double dt, int N_gait, int first, int max_iter, const double* xref,
                               const double* fsteps, double* st_d, int* st_i, double* out, int* info, double* rho,
                               double* dbg, int nthreads) {
  MpcConst C = make_mpc_const(dt, N_gait);
  if (max_iter > 0) C.max_iter = max_iter;
  auto work = [&](int t) {
    std::vector<double> slab(nd::V3_SLAB_DOUBLES, 0.0);
    for (int b = t; b < B; b += nthreads) {
      cta::run_cta(nd::NT, [&]() {
        nd::mpc_v3_solve(C, slab.data(), nd::MpcIo{xref, fsteps, st_d, st_i, out, info, rho, dbg}, b, first, 0u);
      });
    }
    if (nthreads <= 1) g_last_slab = slab;
  };
  if (nthreads <= 1) {
    nthreads = 1;
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
    for (auto& t : th) t.join();
  }
}

// ---- the whole-body tick, groups of G robots per emulated warp (csrc/wbc_tick_core.cuh)
#include "../../quadruped-reactive-walking_b200/csrc/wbc_tick_core.cuh"
extern "C" void emu_wbc_tick_run(int B, int G, int first, const double* q, const double* dq, const double* f_cmd, int f_stride,
                                 const double* contacts, const double* pgoals, const double* vgoals, const double* agoals,
                                 const double* M, double* st, double* f_res, double* ddq_res, int* info, double* ddq_cmd,
                                 double* vdes, double* qdes, double* tau_ff, double* Jc_dbg, double* rnea_dbg) {
  WbcConst C = make_wbc_const();
  double Y[36], Yinv[36], U[36], V[36];
  for (int e = 0; e < 36; ++e) Y[e] = M[(e % 6) + 18 * (e / 6)];
  pinv6_serial(Y, Yinv, U, V);
  wt::TickIo io{q, dq, f_cmd, contacts, pgoals, vgoals, agoals, Yinv, f_stride, st, f_res, ddq_res, info,
                ddq_cmd, vdes, qdes, tau_ff, Jc_dbg, rnea_dbg};
  const int warps = 2;  // two emulated warps per CTA, as the kernel has several
  std::vector<double> slab((size_t)warps * wt::WT2_SLAB_DOUBLES, 0.0);
  int counter = 0;
  cta::run_cta(32 * warps, [&]() {
    const int tid = cta::tid(), lane = tid & 31;
    double* sl = slab.data() + (size_t)(tid >> 5) * wt::WT2_SLAB_DOUBLES;
    for (;;) {
      int b0 = 0;
      if (lane == 0) b0 = cta::fetch_add(&counter, G);
      b0 = (int)cta::shfl((double)b0, 0);
      if (b0 >= B) break;
      wt::tick_group(C, B, G, first, io, sl, lane, b0);
    }
  });
}
