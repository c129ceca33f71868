// tick_abi.cuh -- C ABI (qrw_tick_*, include/qrw_b200.h) of the CLOSED control tick of B robots, entirely in HBM:
//
//     planners (Gait, FootstepPlanner, FootTrajectoryGenerator, StatePlanner, feet commands)      planner_tick_kernel
//     -> every k_mpc-th tick: MPC::run on the planners' xref / fsteps                               mpc_v2_kernel
//     -> whole-body tick with f_cmd = x_f_mpc[12:24, 0], contacts = gait row 0, feet commands       wbc_ctrl_inputs + whole-body tick
//
// = /root/reference/scripts/Controller.py:222-303 (compute()) with MPC_Wrapper's synchronous mode
// (/root/reference/scripts/MPC_Wrapper.py:64-71,122-126: the very first tick uses the default forces [0, 0, 8] per foot,
// every later tick the latest MPC result).  The static-gait branch of Controller.py:266-268 reads attributes that do not
// exist in the reference (`self.planner`) and is not mirrored.
//
// The tick index enters the kernels only through k % k_mpc and k != 0, so the k_mpc ticks of one MPC period are captured
// ONCE into a CUDA graph (qrw_tick_capture_period) and replayed for every later period from per-period staging buffers:
// one graph launch = k_mpc control ticks of the whole batch (qrw_tick_replay_period / qrw_tick_period_host).
// Included at the end of qrw_b200.cu (uses its fail() / CU() helpers and the handle structs).
#pragma once

struct qrw_tick {
  int B = 0, device = 0, k_mpc = 0, n_steps = 0, N_gait = 0;
  double h_ref = 0.0;
  qrw_planner_t* pl = nullptr;
  qrw_mpc_t* mpc = nullptr;
  qrw_wbctick_t* wbc = nullptr;
  double* d_default_f = nullptr;  // [B][12] default contact forces of the first tick
  bool have_result = false;       // MPC_Wrapper.not_first_iter
  // per-period staging: inputs [k_mpc][B][7 | 6 | 6], codes [k_mpc][B]; outputs [k_mpc][B][12]
  double *d_q7 = nullptr, *d_hv = nullptr, *d_vref = nullptr, *d_tau = nullptr, *d_qdes = nullptr, *d_vdes = nullptr;
  int* d_code = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaStream_t stream = nullptr;
};

namespace {
int tick_run(qrw_tick* h, int k, const double* q7, const double* hv, const double* vref, const int* code, cudaStream_t s) {
  if (qrw_planner_run_device(h->pl, k, qrw::plan::ST_ALL, q7, hv, vref, code, s)) return 1;
  const double *xref, *fsteps, *contacts, *p_cmd, *v_cmd, *a_cmd;
  if (qrw_planner_outputs_device(h->pl, &xref, &fsteps, &contacts, nullptr, nullptr, nullptr, nullptr, &p_cmd, &v_cmd, &a_cmd)) return 1;
  if (k % h->k_mpc == 0 && qrw_mpc_run_device(h->mpc, k, xref, fsteps, s)) return 1;  // MPC_Wrapper.solve -> MPC::run(k, xref, fsteps)
  const double* f = h->have_result ? h->mpc->d_out + 12 : h->d_default_f;
  const int stride = h->have_result ? 24 * h->n_steps : 12;
  h->have_result = true;
  return qrw_wbctick_run_ctrl_device(h->wbc, h->h_ref, vref, f, stride, contacts, p_cmd, v_cmd, a_cmd, s);
}
}  // namespace

extern "C" {

int qrw_tick_create(qrw_tick_t** out, int batch, double dt_wbc, double dt_mpc, double T_gait, double T_mpc, int N_gait, double h_ref,
                    const double* q_init12, int device) {
  if (!out) return fail("qrw_tick_create: null output handle");
  if (batch <= 0 || dt_wbc <= 0 || dt_mpc <= 0 || T_mpc <= 0) return fail("qrw_tick_create: bad argument");
  auto* h = new qrw_tick();
  h->B = batch;
  h->device = device;
  h->k_mpc = (int)(dt_mpc / dt_wbc + 0.5);
  h->n_steps = (int)(T_mpc / dt_mpc + 0.5);
  h->N_gait = N_gait;
  h->h_ref = h_ref;
  if (qrw_planner_create(&h->pl, batch, dt_mpc, dt_wbc, T_gait, T_mpc, N_gait, h->k_mpc, h_ref, device) ||
      qrw_mpc_create(&h->mpc, batch, dt_mpc, h->n_steps, T_gait, N_gait, device) || qrw_wbctick_create(&h->wbc, batch, device)) {
    qrw_tick_destroy(h);
    return 1;
  }
  const size_t B = (size_t)batch, K = (size_t)h->k_mpc;
  std::vector<double> qd(B * 12), f0(B * 12);
  static const double q_default[12] = {0.0, 0.7, -1.4, 0.0, 0.7, -1.4, 0.0, -0.7, 1.4, 0.0, -0.7, 1.4};  // main_solo12_control.py q_init
  for (size_t b = 0; b < B; ++b)
    for (int i = 0; i < 12; ++i) {
      qd[b * 12 + i] = q_init12 ? q_init12[i] : q_default[i];
      f0[b * 12 + i] = (i % 3 == 2) ? 8.0 : 0.0;
    }
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaMalloc(&h->d_default_f, B * 12 * sizeof(double));
  if (e == cudaSuccess) e = cudaMemcpy(h->d_default_f, f0.data(), B * 12 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&h->d_q7, K * B * 7 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_hv, K * B * 6 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_vref, K * B * 6 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_code, K * B * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_tau, K * B * 12 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_qdes, K * B * 12 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_vdes, K * B * 12 * sizeof(double));
  if (e == cudaSuccess) e = cudaMemset(h->d_code, 0, K * B * sizeof(int));
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess || qrw_wbctick_set_desired_host(h->wbc, qd.data(), nullptr)) {  // Controller.py:153
    if (e != cudaSuccess) fail(std::string("qrw_tick_create: ") + cudaGetErrorString(e));
    qrw_tick_destroy(h);
    return 1;
  }
  *out = h;
  return 0;
}

int qrw_tick_destroy(qrw_tick_t* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  if (h->exec) cudaGraphExecDestroy(h->exec);
  if (h->graph) cudaGraphDestroy(h->graph);
  qrw_planner_destroy(h->pl);
  qrw_mpc_destroy(h->mpc);
  qrw_wbctick_destroy(h->wbc);
  cudaFree(h->d_default_f), cudaFree(h->d_q7), cudaFree(h->d_hv), cudaFree(h->d_vref), cudaFree(h->d_code);
  cudaFree(h->d_tau), cudaFree(h->d_qdes), cudaFree(h->d_vdes);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

// One control tick k of the whole batch.  Device pointers: q7 [B][7] (x y z, quaternion x y z w), h_v [B][6], v_ref [B][6],
// code [B] joystick codes (NULL = 0).  Asynchronous on `stream`.
int qrw_tick_run_device(qrw_tick_t* h, int k, const double* q7, const double* h_v, const double* v_ref, const int* code, void* stream) {
  if (!h || !q7 || !h_v || !v_ref) return fail("qrw_tick_run_device: null argument");
  if (k < 0) return fail("qrw_tick_run_device: negative tick index");
  CU(cudaSetDevice(h->device));
  return tick_run(h, k, q7, h_v, v_ref, code, (cudaStream_t)stream);
}

// Same with host pointers: H2D of the 156 B per robot, the tick, synchronised on return.
int qrw_tick_run_host(qrw_tick_t* h, int k, const double* q7, const double* h_v, const double* v_ref, const int* code) {
  if (!h || !q7 || !h_v || !v_ref) return fail("qrw_tick_run_host: null argument");
  CU(cudaSetDevice(h->device));
  const size_t B = (size_t)h->B;
  CU(cudaMemcpyAsync(h->d_q7, q7, B * 7 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  CU(cudaMemcpyAsync(h->d_hv, h_v, B * 6 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  CU(cudaMemcpyAsync(h->d_vref, v_ref, B * 6 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  if (code) CU(cudaMemcpyAsync(h->d_code, code, B * sizeof(int), cudaMemcpyHostToDevice, h->stream));
  if (tick_run(h, k, h->d_q7, h->d_hv, h->d_vref, code ? h->d_code : nullptr, h->stream)) return 1;
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}

// Results of the latest tick ([B][12] each; Controller.py:306-310 scales tau_ff by 0.8 on the robot side).
int qrw_tick_result_host(qrw_tick_t* h, double* tau_ff, double* q_des, double* v_des, double* f_res) {
  if (!h) return fail("qrw_tick_result_host: null handle");
  return qrw_wbctick_result_host(h->wbc, tau_ff, q_des, v_des, f_res, nullptr);
}
int qrw_tick_result_device(qrw_tick_t* h, const double** tau_ff, const double** q_des, const double** v_des, const double** f_res) {
  if (!h) return fail("qrw_tick_result_device: null handle");
  return qrw_wbctick_result_device(h->wbc, tau_ff, q_des, v_des, f_res);
}
// The three objects a tick is composed of (state snapshots, getters, debugging).  Any of the outputs may be NULL.
int qrw_tick_parts(qrw_tick_t* h, qrw_planner_t** planner, qrw_mpc_t** mpc, qrw_wbctick_t** wbc) {
  if (!h) return fail("qrw_tick_parts: null handle");
  if (planner) *planner = h->pl;
  if (mpc) *mpc = h->mpc;
  if (wbc) *wbc = h->wbc;
  return 0;
}

// Capture the k_mpc ticks of one MPC period (ticks k_mpc .. 2 k_mpc - 1: warm, the first of them solves the MPC) into a
// CUDA graph reading the per-period staging buffers.  Needs at least one executed tick (the default forces of the very
// first tick are not part of the graph).  Nothing is executed; the handle's state is what it was.
int qrw_tick_capture_period(qrw_tick_t* h) {
  if (!h) return fail("qrw_tick_capture_period: null handle");
  if (!h->have_result) return fail("qrw_tick_capture_period: run the first tick(s) before capturing (the captured ticks are warm)");
  CU(cudaSetDevice(h->device));
  if (h->exec) {
    cudaGraphExecDestroy(h->exec), cudaGraphDestroy(h->graph);
    h->exec = nullptr, h->graph = nullptr;
  }
  CU(cudaDeviceSynchronize());
  const size_t B = (size_t)h->B;
  const double *tau, *qdes, *vdes;
  if (qrw_wbctick_result_device(h->wbc, &tau, &qdes, &vdes, nullptr)) return 1;
  CU(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
  int rc = 0;
  for (int i = 0; i < h->k_mpc && !rc; ++i) {
    rc = tick_run(h, h->k_mpc + i, h->d_q7 + i * B * 7, h->d_hv + i * B * 6, h->d_vref + i * B * 6, h->d_code + i * B, h->stream);
    const size_t n = B * 12 * sizeof(double);
    if (!rc && (cudaMemcpyAsync(h->d_tau + i * B * 12, tau, n, cudaMemcpyDeviceToDevice, h->stream) != cudaSuccess ||
                cudaMemcpyAsync(h->d_qdes + i * B * 12, qdes, n, cudaMemcpyDeviceToDevice, h->stream) != cudaSuccess ||
                cudaMemcpyAsync(h->d_vdes + i * B * 12, vdes, n, cudaMemcpyDeviceToDevice, h->stream) != cudaSuccess))
      rc = fail("qrw_tick_capture_period: copy into the period outputs failed");
  }
  cudaError_t e = cudaStreamEndCapture(h->stream, &h->graph);
  if (rc) return 1;
  if (e != cudaSuccess) return fail(std::string("qrw_tick_capture_period: ") + cudaGetErrorString(e));
  CU(cudaGraphInstantiate(&h->exec, h->graph, 0));
  return 0;
}
// Device staging of one period: inputs q7 [k_mpc][B][7], h_v / v_ref [k_mpc][B][6], code [k_mpc][B]; outputs tau_ff / q_des /
// v_des [k_mpc][B][12].  Any pointer may be NULL.
int qrw_tick_period_buffers(qrw_tick_t* h, double** q7, double** h_v, double** v_ref, int** code, const double** tau_ff,
                            const double** q_des, const double** v_des) {
  if (!h) return fail("qrw_tick_period_buffers: null handle");
  if (q7) *q7 = h->d_q7;
  if (h_v) *h_v = h->d_hv;
  if (v_ref) *v_ref = h->d_vref;
  if (code) *code = h->d_code;
  if (tau_ff) *tau_ff = h->d_tau;
  if (q_des) *q_des = h->d_qdes;
  if (v_des) *v_des = h->d_vdes;
  return 0;
}
// One graph launch = k_mpc control ticks of the batch from the staging buffers.  Asynchronous on `stream`.
int qrw_tick_replay_period(qrw_tick_t* h, void* stream) {
  if (!h) return fail("qrw_tick_replay_period: null handle");
  if (!h->exec) return fail("qrw_tick_replay_period: no captured period (call qrw_tick_capture_period first)");
  CU(cudaSetDevice(h->device));
  CU(cudaGraphLaunch(h->exec, (cudaStream_t)stream));
  ++g_launches;
  return 0;
}
// One MPC period end to end from host memory: H2D of the period's inputs (156 B per robot and tick; code may be NULL = 0),
// one graph launch, D2H of the period's outputs (288 B per robot and tick; any output may be NULL), synchronised on return.
int qrw_tick_period_host(qrw_tick_t* h, const double* q7, const double* h_v, const double* v_ref, const int* code, double* tau_ff,
                         double* q_des, double* v_des) {
  if (!h || !q7 || !h_v || !v_ref) return fail("qrw_tick_period_host: null argument");
  if (!h->exec) return fail("qrw_tick_period_host: no captured period (call qrw_tick_capture_period first)");
  CU(cudaSetDevice(h->device));
  const size_t n = (size_t)h->k_mpc * h->B;
  cudaStream_t s = h->stream;
  CU(cudaMemcpyAsync(h->d_q7, q7, n * 7 * sizeof(double), cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(h->d_hv, h_v, n * 6 * sizeof(double), cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(h->d_vref, v_ref, n * 6 * sizeof(double), cudaMemcpyHostToDevice, s));
  if (code) CU(cudaMemcpyAsync(h->d_code, code, n * sizeof(int), cudaMemcpyHostToDevice, s));
  else CU(cudaMemsetAsync(h->d_code, 0, n * sizeof(int), s));
  CU(cudaGraphLaunch(h->exec, s));
  ++g_launches;
  if (tau_ff) CU(cudaMemcpyAsync(tau_ff, h->d_tau, n * 12 * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (q_des) CU(cudaMemcpyAsync(q_des, h->d_qdes, n * 12 * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (v_des) CU(cudaMemcpyAsync(v_des, h->d_vdes, n * 12 * sizeof(double), cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  return 0;
}

}  // extern "C"
