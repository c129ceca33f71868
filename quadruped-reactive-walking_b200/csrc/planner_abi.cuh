// planner_abi.cuh -- kernel + C ABI (qrw_planner_*, include/qrw_b200.h) of the batched input producers.
// Included at the end of qrw_b200.cu (uses its fail() / CU() / g_launches helpers).
#pragma once
#include <vector>

#include "planner_core.cuh"

namespace {

using qrw::plan::LaneCtx;
using qrw::plan::PlanConst;
using qrw::plan::PlanMem;

constexpr int PLAN_THREADS = 64;  // 16 robots per CTA: the footstep staging (16 x N_gait x 12 doubles) fits 48 KB up to N_gait = 32

// One control tick of the planners of B robots, four lanes per robot (planner_core.cuh).  The footstep matrix of a
// robot is assembled column by column in shared memory and leaves the SM as one contiguous run per warp (8 robots).
__global__ void __launch_bounds__(PLAN_THREADS, 7) planner_tick_kernel(const PlanConst C, const PlanMem M, const int k, const int stages,
                                                                    const double* __restrict__ q7, const double* __restrict__ v6,
                                                                    const double* __restrict__ vref6, const int* __restrict__ code) {
  extern __shared__ __align__(16) double fs_stage[];
  const int t = blockIdx.x * PLAN_THREADS + threadIdx.x;
  const int b = t >> 2, j = t & 3;
  const int fsz = C.N_gait * 12;
  const bool active = b < M.B;
  LaneCtx X;
  if (active) {
    double q[7], hv[6], vr[6];
#pragma unroll
    for (int i = 0; i < 7; ++i) q[i] = q7[(size_t)b * 7 + i];
#pragma unroll
    for (int i = 0; i < 6; ++i) hv[i] = v6[(size_t)b * 6 + i], vr[i] = vref6[(size_t)b * 6 + i];
    qrw::plan::lane_compute(C, M, b, j, k, stages, q, hv, vr, code ? code[b] : 0, fs_stage + (threadIdx.x >> 2) * fsz + 3 * C.N_gait * j, X);
  }
  __syncwarp();
  if (active) qrw::plan::lane_store(C, M, b, j, X);
  if (stages & qrw::plan::ST_FOOTSTEPS) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b0 = (blockIdx.x * PLAN_THREADS >> 2) + 8 * warp;  // first robot of this warp
    int cnt = M.B - b0;
    cnt = cnt > 8 ? 8 : cnt;
    if (cnt > 0) {
      const double2* src = reinterpret_cast<const double2*>(fs_stage + (size_t)8 * warp * fsz);
      double2* dst = reinterpret_cast<double2*>(M.fsteps + (size_t)b0 * fsz);  // fsz is even, the base is 256 B aligned
      const int n2 = cnt * fsz / 2;
      for (int i = lane; i < n2; i += 32) dst[i] = src[i];
    }
  }
}

}  // namespace

struct qrw_planner {
  int B = 0, device = 0;
  PlanConst C;
  PlanMem M;
  double ftg_target[12], ftg_pos[12];
  double *d_q = nullptr, *d_v = nullptr, *d_vref = nullptr;  // staging of the host entry point
  int* d_code = nullptr;
  cudaStream_t stream = nullptr;
};

extern "C" {

// (Re)initialise the state of every robot: Gait::initialize (trot + create_gait_f), FootstepPlanner::initialize,
// FootTrajectoryGenerator::initialize.  The initial gait is the same for the whole batch, so it is built once on the
// host with the kernel's own GaitState code and broadcast.
static int planner_reset(qrw_planner_t* h) {
  const size_t B = (size_t)h->B;
  qrw::plan::GaitState g;
  g.past.clear(), g.cur.clear(), g.des.clear();
  g.newPhase = 0, g.is_static = 1, g.remaining = 0.0;
  g.create(h->C, 3);
  g.create_gait_f(h->C);
  std::vector<unsigned long long> bits(6 * B);
  const qrw::plan::GaitBits* m[3] = {&g.past, &g.cur, &g.des};
  for (int f = 0; f < 3; ++f)
    for (size_t b = 0; b < B; ++b) bits[(f * B + b) * 2] = m[f]->lo, bits[(f * B + b) * 2 + 1] = m[f]->hi;
  std::vector<int> flags(B, 2);  // newPhase = false, is_static = true (Gait.cpp:11-12), no swing feet recorded
  std::vector<double> st((size_t)qrw::plan::PLAN_STATE_DOUBLES * B, 0.0);
  auto field = [&](int f, int len, size_t b) { return st.data() + B * f + b * len; };
  for (size_t b = 0; b < B; ++b) {
    for (int i = 0; i < 12; ++i) {
      field(qrw::plan::F_CURFS, 12, b)[i] = h->C.shoulders[i];    // FootstepPlanner.cpp:38-41
      field(qrw::plan::F_TARGET, 12, b)[i] = h->C.shoulders[i];
      field(qrw::plan::F_OTARGET, 12, b)[i] = h->C.shoulders[i];
      field(qrw::plan::F_FTGTGT, 12, b)[i] = h->ftg_target[i];    // FootTrajectoryGenerator.cpp:35-36
      field(qrw::plan::F_FPOS, 12, b)[i] = h->ftg_pos[i];
    }
    for (int j = 0; j < 4; ++j) field(qrw::plan::F_CONTACTS, 4, b)[j] = g.cur.at(0, j) ? 1.0 : 0.0;
  }
  CU(cudaMemcpy(h->M.bits, bits.data(), bits.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(h->M.flags, flags.data(), B * sizeof(int), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(h->M.st, st.data(), st.size() * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemset(h->M.xref, 0, B * 12 * (h->C.n_steps + 1) * sizeof(double)));
  CU(cudaMemset(h->M.fsteps, 0, B * 12 * h->C.N_gait * sizeof(double)));
  return 0;
}

int qrw_planner_create(qrw_planner_t** out, int batch, double dt_mpc, double dt_wbc, double T_gait, double T_mpc, int N_gait,
                       int k_mpc, double h_ref, int device) {
  if (!out) return fail("qrw_planner_create: null output handle");
  if (batch <= 0 || dt_mpc <= 0 || dt_wbc <= 0 || k_mpc <= 0) return fail("qrw_planner_create: bad argument");
  const int n_steps = (int)lround(T_mpc / dt_mpc);
  if (N_gait > 32 || N_gait % 2) return fail("qrw_planner_create: N_gait must be even and <= 32");
  // Gait.cpp:30-31 throws std::invalid_argument here
  if (n_steps > N_gait || (int)lround(T_gait / dt_mpc) > N_gait)
    return fail("Sizes of matrices are too small for considered durations. Increase N_gait in config file.");
  if (sm_count_of(device) <= 0) return fail("qrw_planner_create: no usable CUDA device (this library has no CPU fallback)");
  CU(cudaSetDevice(device));
  auto* h = new qrw_planner();
  h->B = batch, h->device = device;
  PlanConst& C = h->C;
  C.dt_mpc = dt_mpc, C.dt_wbc = dt_wbc, C.T_gait = T_gait, C.T_mpc = T_mpc, C.h_ref = h_ref;
  C.max_height = 0.05, C.lock_time = 0.07;  // Controller.py:138
  C.lin_step = n_steps > 1 ? (T_mpc - dt_mpc) / (n_steps - 1) : 0.0;
  C.N_gait = N_gait, C.n_steps = n_steps, C.k_mpc = k_mpc, C.pad = 0;
  const double sh[12] = {0.1946, 0.14695, 0.0, 0.1946, -0.14695, 0.0, -0.1946, 0.14695, 0.0, -0.1946, -0.14695, 0.0};  // Controller.py:131-133
  for (int i = 0; i < 12; ++i) C.shoulders[i] = h->ftg_target[i] = h->ftg_pos[i] = sh[i];
  const size_t B = (size_t)batch;
  h->M.B = batch;
  CUH(cudaMalloc(&h->M.st, B * qrw::plan::PLAN_STATE_DOUBLES * sizeof(double)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->M.bits, B * 6 * sizeof(unsigned long long)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->M.flags, B * sizeof(int)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->M.xref, B * 12 * (n_steps + 1) * sizeof(double)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->M.fsteps, B * 12 * N_gait * sizeof(double)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->d_q, B * 7 * sizeof(double)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->d_v, B * 6 * sizeof(double)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->d_vref, B * 6 * sizeof(double)), qrw_planner_destroy(h));
  CUH(cudaMalloc(&h->d_code, B * sizeof(int)), qrw_planner_destroy(h));
  CUH(cudaStreamCreate(&h->stream), qrw_planner_destroy(h));
  if (planner_reset(h)) return 1;
  *out = h;
  return 0;
}

int qrw_planner_destroy(qrw_planner_t* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaFree(h->M.st), cudaFree(h->M.bits), cudaFree(h->M.flags), cudaFree(h->M.xref), cudaFree(h->M.fsteps);
  cudaFree(h->d_q), cudaFree(h->d_v), cudaFree(h->d_vref), cudaFree(h->d_code);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

int qrw_planner_configure(qrw_planner_t* h, const double* shoulders, const double* ftg_target, const double* ftg_position,
                          double max_height, double lock_time) {
  if (!h) return fail("qrw_planner_configure: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  for (int i = 0; i < 12; ++i) {
    if (shoulders) h->C.shoulders[i] = shoulders[i];
    if (ftg_target) h->ftg_target[i] = ftg_target[i];
    if (ftg_position) h->ftg_pos[i] = ftg_position[i];
  }
  h->C.max_height = max_height, h->C.lock_time = lock_time;
  return planner_reset(h);
}

int qrw_planner_run_device(qrw_planner_t* h, int k, int stages, const double* q7, const double* v6, const double* vref6,
                           const int* code, void* stream) {
  if (!h || !q7 || !v6 || !vref6) return fail("qrw_planner_run_device: null argument");
  if (k < 0 || (stages & ~qrw::plan::ST_ALL)) return fail("qrw_planner_run_device: bad tick index or stage mask");
  CU(cudaSetDevice(h->device));
  const unsigned grid = (unsigned)((4LL * h->B + PLAN_THREADS - 1) / PLAN_THREADS);
  const size_t smem = (size_t)(PLAN_THREADS / 4) * h->C.N_gait * 12 * sizeof(double);
  planner_tick_kernel<<<grid, PLAN_THREADS, smem, (cudaStream_t)stream>>>(h->C, h->M, k, stages, q7, v6, vref6, code);
  ++g_launches;
  CU(cudaGetLastError());
  return 0;
}

int qrw_planner_run_host(qrw_planner_t* h, int k, int stages, const double* q7, const double* v6, const double* vref6,
                         const int* code) {
  if (!h || !q7 || !v6 || !vref6) return fail("qrw_planner_run_host: null argument");
  CU(cudaSetDevice(h->device));
  const size_t B = (size_t)h->B;
  cudaStream_t s = h->stream;
  CU(cudaMemcpyAsync(h->d_q, q7, B * 7 * sizeof(double), cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(h->d_v, v6, B * 6 * sizeof(double), cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(h->d_vref, vref6, B * 6 * sizeof(double), cudaMemcpyHostToDevice, s));
  if (code) CU(cudaMemcpyAsync(h->d_code, code, B * sizeof(int), cudaMemcpyHostToDevice, s));
  if (qrw_planner_run_device(h, k, stages, h->d_q, h->d_v, h->d_vref, code ? h->d_code : nullptr, s)) return 1;
  CU(cudaStreamSynchronize(s));
  return 0;
}

static void planner_unpack_gait(const unsigned long long* w, int N, double* out) {  // N x 4 column-major doubles
  qrw::plan::GaitBits g;
  g.lo = w[0], g.hi = w[1];
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < 4; ++j) out[i + N * j] = g.at(i, j) ? 1.0 : 0.0;
}

// which = 0 past, 1 current, 2 desired gait -> [B][N_gait*4]
static int planner_gait_host(qrw_planner_t* h, int which, double* out) {
  const size_t B = (size_t)h->B;
  std::vector<unsigned long long> w(2 * B);
  CU(cudaMemcpy(w.data(), h->M.bits + (size_t)which * B * 2, 2 * B * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  for (size_t b = 0; b < B; ++b) planner_unpack_gait(&w[2 * b], h->C.N_gait, out + b * h->C.N_gait * 4);
  return 0;
}

int qrw_planner_outputs_device(qrw_planner_t* h, const double** xref, const double** fsteps, const double** contacts,
                               const double** o_target, const double** fpos, const double** fvel, const double** facc,
                               const double** p_cmd, const double** v_cmd, const double** a_cmd) {
  if (!h) return fail("qrw_planner_outputs_device: null handle");
  const size_t B = (size_t)h->B;
  if (xref) *xref = h->M.xref;
  if (fsteps) *fsteps = h->M.fsteps;
  if (contacts) *contacts = h->M.st + B * qrw::plan::F_CONTACTS;
  if (o_target) *o_target = h->M.st + B * qrw::plan::F_OTARGET;
  if (fpos) *fpos = h->M.st + B * qrw::plan::F_FPOS;
  if (fvel) *fvel = h->M.st + B * qrw::plan::F_FVEL;
  if (facc) *facc = h->M.st + B * qrw::plan::F_FACC;
  if (p_cmd) *p_cmd = h->M.st + B * qrw::plan::F_PCMD;
  if (v_cmd) *v_cmd = h->M.st + B * qrw::plan::F_VCMD;
  if (a_cmd) *a_cmd = h->M.st + B * qrw::plan::F_ACMD;
  return 0;
}

int qrw_planner_outputs_host(qrw_planner_t* h, double* xref, double* fsteps, double* gait, double* o_target, double* fpos,
                             double* fvel, double* facc, double* p_cmd, double* v_cmd, double* a_cmd) {
  if (!h) return fail("qrw_planner_outputs_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  const size_t B = (size_t)h->B;
  if (xref) CU(cudaMemcpy(xref, h->M.xref, B * 12 * (h->C.n_steps + 1) * sizeof(double), cudaMemcpyDeviceToHost));
  if (fsteps) CU(cudaMemcpy(fsteps, h->M.fsteps, B * 12 * h->C.N_gait * sizeof(double), cudaMemcpyDeviceToHost));
  if (gait && planner_gait_host(h, 1, gait)) return 1;
  double* dst[7] = {o_target, fpos, fvel, facc, p_cmd, v_cmd, a_cmd};
  const int fld[7] = {qrw::plan::F_OTARGET, qrw::plan::F_FPOS, qrw::plan::F_FVEL, qrw::plan::F_FACC, qrw::plan::F_PCMD,
                      qrw::plan::F_VCMD, qrw::plan::F_ACMD};
  for (int i = 0; i < 7; ++i)
    if (dst[i]) CU(cudaMemcpy(dst[i], h->M.st + B * fld[i], B * 12 * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

int qrw_planner_state_host(qrw_planner_t* h, double* past_gait, double* desired_gait, double* current_footstep, double* t0s,
                           double* t_swing, int* flags) {
  if (!h) return fail("qrw_planner_state_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  const size_t B = (size_t)h->B;
  if (past_gait && planner_gait_host(h, 0, past_gait)) return 1;
  if (desired_gait && planner_gait_host(h, 2, desired_gait)) return 1;
  if (current_footstep)
    CU(cudaMemcpy(current_footstep, h->M.st + B * qrw::plan::F_CURFS, B * 12 * sizeof(double), cudaMemcpyDeviceToHost));
  if (t0s) CU(cudaMemcpy(t0s, h->M.st + B * qrw::plan::F_T0S, B * 4 * sizeof(double), cudaMemcpyDeviceToHost));
  if (t_swing) CU(cudaMemcpy(t_swing, h->M.st + B * qrw::plan::F_TSW, B * 4 * sizeof(double), cudaMemcpyDeviceToHost));
  if (flags) {
    std::vector<int> f(B);
    CU(cudaMemcpy(f.data(), h->M.flags, B * sizeof(int), cudaMemcpyDeviceToHost));
    for (size_t b = 0; b < B; ++b) flags[2 * b] = f[b] & 1, flags[2 * b + 1] = (f[b] >> 1) & 1;
  }
  return 0;
}

// FootTrajectoryGenerator::update takes its target as an argument (FootTrajectoryGenerator.cpp:110); the fused tick
// feeds it FootstepPlanner's o_targetFootstep_.  This overwrites that field for callers that drive the stages one by one.
int qrw_planner_set_target_host(qrw_planner_t* h, const double* o_target) {
  if (!h || !o_target) return fail("qrw_planner_set_target_host: null argument");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(h->M.st + (size_t)h->B * qrw::plan::F_OTARGET, o_target, (size_t)h->B * 12 * sizeof(double), cudaMemcpyHostToDevice));
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------- control-loop glue (Controller.py:258-303)
namespace {
// q_wbc = (0, 0, h_ref, 0, 0, 0, 1, qdes of the previous tick), b_v = (v_ref[0:6], vdes of the previous tick)
// (Controller.py:275-283) and f_cmd = x_f_mpc[12:24, 0] (:258, :300) gathered from the MPC result with its stride.
// Twelve lanes per robot, one per joint / force component.
__global__ void __launch_bounds__(192) wbc_ctrl_inputs_kernel(const int B, const double h_ref, const double* __restrict__ vref6,
                                                             const double* __restrict__ f_src, const int f_stride,
                                                             const double* __restrict__ qdes, const double* __restrict__ vdes,
                                                             double* __restrict__ q, double* __restrict__ dq, double* __restrict__ f) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t b = t / 12;
  const int i = (int)(t % 12);
  if (b >= (size_t)B) return;
  q[b * 19 + 7 + i] = qdes[b * 12 + i];
  dq[b * 18 + 6 + i] = vdes[b * 12 + i];
  f[b * 12 + i] = f_src[b * (size_t)f_stride + i];
  if (i < 7) q[b * 19 + i] = (i == 2) ? h_ref : (i == 6 ? 1.0 : 0.0);
  if (i < 6) dq[b * 18 + i] = vref6[b * 6 + i];
}
}  // namespace

extern "C" {

// wbc_controller.qdes[7:] / vdes[6:] before the first tick (Controller.py:153: q_init, zeros).  Host pointers [B][12];
// NULL = zeros.
int qrw_wbctick_set_desired_host(qrw_wbctick_t* h, const double* qdes, const double* vdes) {
  if (!h) return fail("qrw_wbctick_set_desired_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  const size_t n = (size_t)h->B * 12 * sizeof(double);
  if (qdes) CU(cudaMemcpy(h->d_qdes, qdes, n, cudaMemcpyHostToDevice));
  else CU(cudaMemset(h->d_qdes, 0, n));
  if (vdes) CU(cudaMemcpy(h->d_vdes, vdes, n, cudaMemcpyHostToDevice));
  else CU(cudaMemset(h->d_vdes, 0, n));
  return 0;
}

// The whole-body tick as scripts/Controller.py:275-303 calls it inside the control loop: the configuration and velocity
// handed to wbc_controller.compute are built from the PREVIOUS tick's qdes / vdes held by the handle, the reference
// velocity and the reference height; f_cmd is read with a stride (f_cmd + b * f_stride, e.g. the MPC result + 12 with
// stride 24 * n_steps).  Device pointers, asynchronous on `stream`.
int qrw_wbctick_run_ctrl_device(qrw_wbctick_t* h, double h_ref, const double* vref6, const double* f_cmd, int f_stride,
                                const double* contacts, const double* pgoals, const double* vgoals, const double* agoals,
                                void* stream) {
  if (!h || !vref6 || !f_cmd || !contacts || !pgoals || !vgoals || !agoals) return fail("qrw_wbctick_run_ctrl_device: null argument");
  if (f_stride < 12) return fail("qrw_wbctick_run_ctrl_device: f_stride < 12");
  CU(cudaSetDevice(h->device));
  const unsigned grid = (unsigned)((12LL * h->B + 191) / 192);
  wbc_ctrl_inputs_kernel<<<grid, 192, 0, (cudaStream_t)stream>>>(h->B, h_ref, vref6, f_cmd, f_stride, h->d_qdes, h->d_vdes, h->d_q,
                                                                h->d_dq, h->d_f);
  ++g_launches;
  CU(cudaGetLastError());
  return qrw_wbctick_run_device(h, h->d_q, h->d_dq, h->d_f, contacts, pgoals, vgoals, agoals, stream);
}

}  // extern "C"
