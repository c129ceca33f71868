// noalloc_abi.cuh -- ONE whole-body tick without any allocation: qrw_wbctick_noalloc_device.  Included at the end of qrw_b200.cu.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it); every other entry point of this library works on a
// handle whose buffers qrw_*_create allocates.  This one runs the same wbc_tick_kernel (csrc/wbc_tick_core.cuh) with the
// kernel's scratch -- the QP's state / ddq_res / info / ddq_cmd per robot, the work counter of the persistent grid and
// pinv(M[:6, :6]) -- in a file-scope __device__ array of fixed size, so it calls no cudaMalloc / cudaFree / allocating
// routine at all.  The caller passes inputs and outputs (device pointers).
//
// LIMITS AND RULES
//   * at most QRW_WBCTICK_NOALLOC_MAX_ROBOTS (4096, include/qrw_b200.h) robots per call; a larger batch is rejected with an error;
//   * the array (3.0 MB) is held by every GPU that loads the library, until the process exits;
//   * every call on a device shares the array: calls that overlap on the GPU (two streams, two threads) RACE -- serialise them;
//   * the array keeps the last call's contents.  What must start at zero -- the work counter -- is cleared by every call
//     (cudaMemsetAsync on the caller's stream); the per-robot scratch is written before it is read; pinv(M) is written once, at
//     set-up, and never again;
//   * the QP is always cold-started (QPWBC not yet initialised, src/QPWBC.cpp:23,254): a call equals the FIRST
//     qrw_wbctick_run_device of a fresh handle.  The warm start from the previous tick needs state that survives between calls,
//     which is exactly what a handle is for;
//   * the first call on a device does the set-up (symbol address, constants, one synchronisation): make it outside graph capture.
#pragma once

#include <mutex>

namespace {

constexpr int QRW_NOALLOC_MAX = QRW_WBCTICK_NOALLOC_MAX_ROBOTS;  // robots per call (include/qrw_b200.h)
// layout of the scratch array (doubles)
constexpr int NA_YINV = 0;                                             // [36]   pinv(M[:6, :6]), written at set-up
constexpr int NA_M = 64;                                               // [324]  M (crba at the zero configuration), set-up only
constexpr int NA_COUNTER = 448;                                        // [1 int] work counter of the persistent grid (+ pad)
constexpr int NA_STATE = 512;                                          // [MAX][72] QP state (written; never read: cold start)
constexpr int NA_DDQRES = NA_STATE + QRW_NOALLOC_MAX * WBC_STATE_DOUBLES;  // [MAX][6]
constexpr int NA_DDQCMD = NA_DDQRES + QRW_NOALLOC_MAX * 6;                 // [MAX][12]
constexpr int NA_INFO = NA_DDQCMD + QRW_NOALLOC_MAX * 12;                  // [MAX][4] ints
constexpr int NA_DOUBLES = NA_INFO + QRW_NOALLOC_MAX * 2;

// The scratch itself.  In HOST code this name is a host-side stand-in, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress (noalloc_setup below), once per device.
static __device__ __align__(256) double g_noalloc_scratch[NA_DOUBLES];

constexpr int NA_MAX_DEVICES = 64;
struct NoallocDevice {
  double* base = nullptr;  // device address of g_noalloc_scratch on this device
  int sms = 0;
};
NoallocDevice g_noalloc_dev[NA_MAX_DEVICES];
std::mutex g_noalloc_mutex;

// once per device, with that device current
int noalloc_setup(int device, NoallocDevice** out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  NoallocDevice& d = g_noalloc_dev[device];
  if (!d.base) {
    double* p = nullptr;
    CU(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    if (!p) return fail("qrw_wbctick_noalloc_device: cudaGetSymbolAddress returned a null address");
    const int sms = sm_count_of(device);
    if (sms <= 0) return fail("qrw_wbctick_noalloc_device: cannot read the SM count of the device");
    double M[324] = {0.0}, Y[6];
    solo12_base_inertia(Y);
    for (int i = 0; i < 6; ++i) M[i + 18 * i] = Y[i];
    CU(cudaMemcpy(p + NA_M, M, sizeof(M), cudaMemcpyHostToDevice));
    wbc_yinv_kernel<<<1, 32>>>(p + NA_M, p + NA_YINV);
    ++g_launches;
    CU(cudaGetLastError());
    CU(cudaFuncSetAttribute(wbc_tick_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WT_WARPS * wt::WT2_SLAB_BYTES));
    CU(cudaDeviceSynchronize());
    d.sms = sms;
    d.base = p;
  }
  *out = &d;
  return 0;
}

}  // namespace

extern "C" {

int qrw_wbctick_noalloc_device(int batch, int device, const double* q, const double* dq, const double* f_cmd, const double* contacts,
                               const double* pgoals, const double* vgoals, const double* agoals, double* tau_ff, double* qdes,
                               double* vdes, double* f_res, void* stream) {
  if (!q || !dq || !f_cmd || !contacts || !pgoals || !vgoals || !agoals || !tau_ff || !qdes || !vdes || !f_res)
    return fail("qrw_wbctick_noalloc_device: null argument");
  if (batch <= 0) return fail("qrw_wbctick_noalloc_device: batch must be positive");
  if (batch > QRW_NOALLOC_MAX)
    return fail("qrw_wbctick_noalloc_device: batch " + std::to_string(batch) + " exceeds the fixed scratch of " +
                std::to_string(QRW_NOALLOC_MAX) + " robots (use a qrw_wbctick_create handle, or split the batch into serialised calls)");
  if (device < 0 || device >= NA_MAX_DEVICES) return fail("qrw_wbctick_noalloc_device: device index out of range");
  CU(cudaSetDevice(device));
  NoallocDevice* d = nullptr;
  if (noalloc_setup(device, &d)) return 1;
  cudaStream_t s = (cudaStream_t)stream;
  double* sc = d->base;
  int* counter = reinterpret_cast<int*>(sc + NA_COUNTER);
  CU(cudaMemsetAsync(counter, 0, sizeof(int), s));  // the one thing that must start at zero
  const int slots = WT_CTAS_PER_SM * d->sms * WT_WARPS;
  const int G = batch >= 8 * slots ? 8 : (batch >= 4 * slots ? 4 : 2);  // as qrw_wbctick_run_device chooses it
  int grid = WT_CTAS_PER_SM * d->sms;
  const int need = ((batch + G - 1) / G + WT_WARPS - 1) / WT_WARPS;
  if (grid > need) grid = need;
  wt::TickIo io{q, dq, f_cmd, contacts, pgoals, vgoals, agoals, sc + NA_YINV, 12, sc + NA_STATE, f_res, sc + NA_DDQRES,
                reinterpret_cast<int*>(sc + NA_INFO), sc + NA_DDQCMD, vdes, qdes, tau_ff, nullptr, nullptr};
  wbc_tick_kernel<<<grid, 32 * WT_WARPS, WT_WARPS * wt::WT2_SLAB_BYTES, s>>>(make_wbc_const(), batch, G, /*first=*/1, io, counter);
  ++g_launches;
  CU(cudaGetLastError());
  return 0;
}

}  // extern "C"
