// qrw_b200.cu -- kernels + C ABI (include/qrw_b200.h) of the batched controller core.
//
// Build (see __graft_entry__.build()):
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared \
//        -Xcompiler -fPIC -o libqrw_b200.so qrw_b200.cu
// Plain CUDA runtime only -- no torch, no cuBLAS.  There is no CPU fallback in this library.
//
// Execution model of the MPC (mpc_v2_kernel): a persistent grid of 256-thread CTAs, two per SM (128 registers, 89 KB of
// shared memory and 256 tensor-memory columns each), one robot per CTA at a time; CTAs pull robot indices from a global
// work counter, so the large spread of ADMM iteration counts across a batch (25 ... 4000) does not leave SMs idle
// behind a slow robot.  The shipped library contains this ONE MPC kernel; the two round-1 kernels (warp per robot, 24x24
// block CTA) and the nested-dissection experiment (mpc_v3_core.cuh) are compiled only with -DQRW_AB_KERNELS into the A/B
// test library (qrw_mpc_set_kernel selects among them there).
#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/qrw_b200.h"
#include "mpc_v2_core.cuh"
#ifdef QRW_AB_KERNELS  // A/B-test build only (tests/ab/libqrw_b200_ab.so): the round-1 kernels and the nested-dissection experiment
#include "mpc_core.cuh"
#include "mpc_cta_core.cuh"
#include "mpc_v3_core.cuh"
#include "mpc_v4_core.cuh"
#endif
#include "wbc_core.cuh"
#include "solo12_core.cuh"
#include "wbc_tick_core.cuh"

using namespace qrw;

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};  // incremented from any calling thread

int fail(const std::string& what) {
  g_err = what;
  return 1;
}
#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail(std::string(#call) + ": " + cudaGetErrorString(e_));                    \
  } while (0)
// inside a *_create after the handle exists: a failed allocation releases everything allocated so far (the destroy functions
// tolerate null members) instead of leaking the handle and its device / pinned memory
#define CUH(call, destroy)                                                                \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      fail(std::string(#call) + ": " + cudaGetErrorString(e_));                           \
      destroy;                                                                            \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)

// ------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------
#ifdef QRW_AB_KERNELS
__global__ void __launch_bounds__(32) mpc_solve_kernel(const MpcConst C, const int B, const int first,
                                                       const double* __restrict__ xref,
                                                       const double* __restrict__ fsteps, double* st_d,
                                                       int* st_i, double* out, int* info, double* rho,
                                                       double* dbg, int* counter) {
  extern __shared__ __align__(16) double slab[];
  const int lane = threadIdx.x;
  for (;;) {
    int b = 0;
    if (lane == 0) b = atomicAdd(counter, 1);
    b = __shfl_sync(0xffffffffu, b, 0);
    if (b >= B) break;
    mpc_solve_instance(C, slab, xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * C.N_gait,
                       st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, first,
                       out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr);
    __syncwarp();
  }
}

// CTA-per-instance MPC solve (mpc_cta_core.cuh): persistent grid of 256-thread CTAs, two per SM, pulling
// instance indices from the global work counter.
__global__ void __launch_bounds__(cta::NT, 2) mpc_cta_kernel(const MpcConst C, const int B, const int first,
                                                            const double* __restrict__ xref,
                                                            const double* __restrict__ fsteps, double* st_d,
                                                            int* st_i, double* out, int* info, double* rho,
                                                            double* dbg, int* counter) {
  extern __shared__ __align__(16) double slab[];
  int* next = reinterpret_cast<int*>(slab + cta::C_INT) + 127;
  for (;;) {
    if (threadIdx.x == 0) *next = atomicAdd(counter, 1);
    __syncthreads();
    const int b = *next;
    if (b >= B) break;
    cta::mpc_cta_solve(C, slab, xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * C.N_gait,
                       st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, first,
                       out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr);
    __syncthreads();
  }
}

#endif  // QRW_AB_KERNELS

// Force-first CTA-per-instance MPC solve (mpc_v2_core.cuh): persistent grid of 256-thread CTAs, two per SM
// (128 registers per thread: at 80 the per-thread solver state spilled to local memory and the long-scoreboard
// stalls of those reloads dominated, profiles/r01_ncu_mpc_v2*.txt), pulling instance indices from the global
// work counter.  The
// warp that runs the sequential state recurrences differs between the CTAs that share an SM, so that they
// do not pile up on one SM sub-partition.
__global__ void __launch_bounds__(v2::NT, 2) mpc_v2_kernel(const MpcConst C, const int B, const int first,
                                                           const int sms, const double* __restrict__ xref,
                                                           const double* __restrict__ fsteps, double* st_d,
                                                           int* st_i, double* out, int* info, double* rho,
                                                           double* dbg, int* counter, const int* __restrict__ k_arr,
                                                           const int* __restrict__ order, double* __restrict__ out_host) {
  extern __shared__ __align__(16) double slab[];
  int* next = reinterpret_cast<int*>(slab + v2::V_INT) + 127;
  // first of the three chain warps (elimination, helper, back-substitution): warps 0..2 in the first wave of CTAs, 2..4 in the
  // second, so that the two CTAs that share an SM keep their elimination warps on different SM sub-partitions and the second
  // CTA's elimination shares one with the first CTA's back-substitution.  Measured (tools/ab_probe.py, 4096 robots): offset 2
  // (or 6) 162.1 k solves/s, offset 1 (round 1's choice) 160.4 k, offsets 0 / 4 160.5 k, offset 3 159.5 k.
  const int recw = ((blockIdx.x / sms) * 2) & 7;
  // 256 of the SM's 512 tensor-memory columns for the whole life of this (persistent) CTA: the matrix rows of the solve
  const unsigned tmem = cta::tmem_alloc(reinterpret_cast<unsigned*>(slab + v2::V_INT) + v2::V2_TMEM_SLOT);
  for (;;) {
    if (threadIdx.x == 0) {
      const int idx = atomicAdd(counter, 1);
      *next = idx < B ? (order ? order[idx] : idx) : B;  // longest-expected solves first (mpc_order_kernel)
    }
    __syncthreads();
    const int b = *next;
    if (b >= B) break;
    int f = first;
    if (k_arr) {  // per-instance call index (qrw_mpc_run_indexed_device): < 0 = no call for this robot in this launch
      const int kb = k_arr[b];
      if (kb < 0) {
        __syncthreads();
        continue;
      }
      f = kb == 0 ? 1 : 0;
    }
    v2::mpc_v2_solve(C, slab, xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * C.N_gait,
                     st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, f,
                     out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr, recw, tmem);
    __syncthreads();
    if (out_host) {  // host entry point: the robot's result also goes straight to the pinned (UVA-mapped) host buffer, under the
                     // solves of the other robots, instead of one device-to-host copy of everything behind the launch
      const double* o = out + (size_t)b * 24 * NS;
      double* oh = out_host + (size_t)b * 24 * NS;
      for (int i = threadIdx.x; i < 24 * NS; i += v2::NT) oh[i] = o[i];
    }
  }
  cta::tmem_free(tmem);
}

#ifdef QRW_AB_KERNELS
// Super-stage chain variant (mpc_v4_core.cuh): same launch shape as mpc_v2_kernel; four chain warps instead of three
// Force-first CTA-per-instance MPC solve (mpc_v4_core.cuh): persistent grid of 256-thread CTAs, two per SM
// (128 registers per thread: at 80 the per-thread solver state spilled to local memory and the long-scoreboard
// stalls of those reloads dominated, profiles/r01_ncu_mpc_v2*.txt), pulling instance indices from the global
// work counter.  The
// warp that runs the sequential state recurrences differs between the CTAs that share an SM, so that they
// do not pile up on one SM sub-partition.
__global__ void __launch_bounds__(v4::NT, 2) mpc_v4_kernel(const MpcConst C, const int B, const int first,
                                                           const int sms, const double* __restrict__ xref,
                                                           const double* __restrict__ fsteps, double* st_d,
                                                           int* st_i, double* out, int* info, double* rho,
                                                           double* dbg, int* counter, const int* __restrict__ k_arr,
                                                           const int* __restrict__ order) {
  extern __shared__ __align__(16) double slab[];
  int* next = reinterpret_cast<int*>(slab + v4::V_INT) + 127;
  const int recw = (blockIdx.x / sms) & 3;
  // 256 of the SM's 512 tensor-memory columns for the whole life of this (persistent) CTA: the matrix rows of the solve
  const unsigned tmem = cta::tmem_alloc(reinterpret_cast<unsigned*>(slab + v4::V_INT) + v4::V4_TMEM_SLOT);
  for (;;) {
    if (threadIdx.x == 0) {
      const int idx = atomicAdd(counter, 1);
      *next = idx < B ? (order ? order[idx] : idx) : B;  // longest-expected solves first (mpc_order_kernel)
    }
    __syncthreads();
    const int b = *next;
    if (b >= B) break;
    int f = first;
    if (k_arr) {  // per-instance call index (qrw_mpc_run_indexed_device): < 0 = no call for this robot in this launch
      const int kb = k_arr[b];
      if (kb < 0) {
        __syncthreads();
        continue;
      }
      f = kb == 0 ? 1 : 0;
    }
    v4::mpc_v4_solve(C, slab, xref + (size_t)b * 12 * (NS + 1), fsteps + (size_t)b * 12 * C.N_gait,
                     st_d + (size_t)b * MPC_STATE_DOUBLES, st_i + (size_t)b * MPC_STATE_INTS, f,
                     out + (size_t)b * 24 * NS, info + 4 * b, rho + b, dbg ? dbg + (size_t)b * 1281 : nullptr, recw, tmem);
    __syncthreads();
  }
  cta::tmem_free(tmem);
}
#endif  // QRW_AB_KERNELS

#ifdef QRW_AB_KERNELS
// Nested-dissection CTA-per-instance MPC solve (mpc_v3_core.cuh, the default): persistent grid of 256-thread CTAs, two per
// SM (128 registers per thread, 113 KB of shared memory and 256 tensor-memory columns each), pulling instance indices
// from the global work counter.
__global__ void __launch_bounds__(nd::NT, 2) mpc_v3_kernel(const MpcConst C, const int B, const int first,
                                                           const double* __restrict__ xref,
                                                           const double* __restrict__ fsteps, double* st_d,
                                                           int* st_i, double* out, int* info, double* rho,
                                                           double* dbg, int* counter, const int* __restrict__ k_arr,
                                                           const int* __restrict__ order) {
  extern __shared__ __align__(16) double slab[];
  int* next = reinterpret_cast<int*>(slab + nd::V_INT) + 127;
  // 256 of the SM's 512 tensor-memory columns for the whole life of this (persistent) CTA: the matrix rows of the solve
  const unsigned tmem = cta::tmem_alloc(reinterpret_cast<unsigned*>(slab + nd::V_INT) + nd::V3_TMEM_SLOT);
  for (;;) {
    if (threadIdx.x == 0) {
      const int idx = atomicAdd(counter, 1);
      *next = idx < B ? (order ? order[idx] : idx) : B;  // longest-expected solves first (mpc_order_kernel)
    }
    __syncthreads();
    const int b = *next;
    if (b >= B) break;
    int f = first;
    if (k_arr) {  // per-instance call index (qrw_mpc_run_indexed_device): < 0 = no call for this robot in this launch
      const int kb = k_arr[b];
      if (kb < 0) {
        __syncthreads();
        continue;
      }
      f = kb == 0 ? 1 : 0;
    }
    nd::mpc_v3_solve(C, slab, nd::MpcIo{xref, fsteps, st_d, st_i, out, info, rho, dbg}, b, f, tmem);
    __syncthreads();
  }
  cta::tmem_free(tmem);
}
#endif  // QRW_AB_KERNELS

// Work order of a launch: robots sorted by the EXPECTED length of their solve, longest first (counting sort, one CTA;
// iteration counts are multiples of check_every = 25 up to max_iter).  The persistent grid pulls robots in this order, so the
// long solves start early and the grid drains evenly; each robot's solve is unchanged -- only when it starts.
// Why: on config 2 (4096 robots over 296 CTA slots) the iteration counts spread from 250 to 2,500 (mean 580), and in index
// order a long solve that starts late leaves the other 295 slots idle: a list-scheduling model on the measured counts
// (tools/tail_model.py, profiles/r02u_tail_model.txt) puts the launch 14 % over the ideal, 0.7 % with the true counts known.
// The previous tick's count alone does NOT work (measured twice: no gain; the model says why: 15..50 robots per tick jump from
// ~300 to 1,500+ iterations, and one of them at the END of a longest-first queue costs its whole solve).  Those jumps happen to
// robots that are commanded to move fast (friction cones active), which the inputs show: the expectation used is
//   max(previous iteration count, speed_gain * |v_ref,xy|)   with v_ref = xref[6:8, 1]  (speed_gain in iterations per m/s).
__global__ void __launch_bounds__(1024) mpc_order_kernel(const int* __restrict__ info, const double* __restrict__ xref, const double speed_gain,
                                                         const int B, int* __restrict__ order) {
  constexpr int NB = 256;
  __shared__ int hist[NB], base[NB];
  for (int i = threadIdx.x; i < NB; i += blockDim.x) hist[i] = 0;
  __syncthreads();
  auto bucket = [&](int b) {
    const double* x = xref + (size_t)b * 12 * (NS + 1) + 12;  // column 1: the first reference state
    const double sp = speed_gain * sqrt(x[6] * x[6] + x[7] * x[7]);
    const int it = info[4 * b];
    const int e = sp > (double)it ? (sp < 1e6 ? (int)sp : 1000000) : it;  // a NaN speed compares false: the count is used
    int q = NB - 1 - e / 25;  // bucket 0 = the longest
    return q < 0 ? 0 : q;
  };
  for (int b = threadIdx.x; b < B; b += blockDim.x) atomicAdd(&hist[bucket(b)], 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int i = 0; i < NB; ++i) base[i] = acc, acc += hist[i];
  }
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += blockDim.x) order[atomicAdd(&base[bucket(b)], 1)] = b;
}

// gait / S_gait as the doubles the reference getters return (MPC.cpp:770-780)
__global__ void mpc_unpack_gait_kernel(const int* st_i, int B, int N_gait, double* gait, double* sgait) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int per = 4 * N_gait + 12 * NS;
  if (t >= B * per) return;
  const int b = t / per, e = t - b * per;
  const int* s = st_i + (size_t)b * MPC_STATE_INTS;
  if (e < 4 * N_gait) {
    const int i = e / N_gait, k = e - N_gait * i;  // column-major N_gait x 4
    gait[(size_t)b * 4 * N_gait + e] = (double)((s[k] >> i) & 1);
  } else {
    const int q = e - 4 * N_gait, k = q / 12, a = q - 12 * k;
    sgait[(size_t)b * 12 * NS + q] = (double)((s[32 + k] >> (a / 3)) & 1);
  }
}

constexpr int WBC_WARPS = 4;
__global__ void __launch_bounds__(32 * WBC_WARPS) wbc_solve_kernel(const WbcConst C, const int B, const int first,
                                                                  const double* __restrict__ M,
                                                                  const double* __restrict__ Jc,
                                                                  const double* __restrict__ f_cmd,
                                                                  const double* __restrict__ RNEA, double* st,
                                                                  double* f_res, double* ddq_res, double* H,
                                                                  int* info, int* counter, const int m_stride,
                                                                  const int f_stride) {
  extern __shared__ __align__(16) double slab[];
  double* sl = slab + (threadIdx.x >> 5) * WBC_SLAB_DOUBLES;
  const int lane = threadIdx.x & 31;
  for (;;) {
    int b = 0;
    if (lane == 0) b = atomicAdd(counter, 1);
    b = __shfl_sync(0xffffffffu, b, 0);
    if (b >= B) break;
    wbc_solve_instance(C, sl, M + (size_t)b * m_stride, Jc + (size_t)b * 216, f_cmd + (size_t)b * f_stride,
                       RNEA + (size_t)b * 6, st + (size_t)b * WBC_STATE_DOUBLES, first, f_res + (size_t)b * 12,
                       ddq_res + (size_t)b * 6, H ? H + (size_t)b * 144 : nullptr, info + 4 * b);
    __syncwarp();
  }
}

// InvKin: four lanes per instance (one per foot), eight instances per warp
__global__ void __launch_bounds__(128) invkin_kernel(const int B, const double* __restrict__ contacts,
                                                     const double* __restrict__ goals,
                                                     const double* __restrict__ vgoals,
                                                     const double* __restrict__ agoals,
                                                     const double* __restrict__ posf, const double* __restrict__ vf,
                                                     const double* __restrict__ wf, const double* __restrict__ af,
                                                     const double* __restrict__ Jf, double* ddq, double* dq_cmd,
                                                     double* q_step) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = t >> 2, i = t & 3;
  if (b >= B) return;
  const double Kp = 100.0, Kd = 20.0;
  double e[3], a[3], v[3], w[3], vg[3];
  const size_t o12 = (size_t)b * 12;
  for (int c = 0; c < 3; ++c) {
    v[c] = vf[o12 + i + 4 * c];
    w[c] = wf[o12 + i + 4 * c];
    vg[c] = vgoals[o12 + c + 3 * i];
    e[c] = goals[o12 + c + 3 * i] - posf[o12 + i + 4 * c];
    a[c] = Kp * e[c] - Kd * (v[c] - vg[c]) + agoals[o12 + c + 3 * i];
  }
  if (contacts[(size_t)b * 4 + i] != 0.0)
    for (int c = 0; c < 3; ++c) a[c] *= 0.0;
  const double cr[3] = {w[1] * v[2] - w[2] * v[1], w[2] * v[0] - w[0] * v[2], w[0] * v[1] - w[1] * v[0]};
  for (int c = 0; c < 3; ++c) a[c] -= af[o12 + i + 4 * c] + cr[c];
  const double* J = Jf + (size_t)b * 144;
  double Mm[3][3], I[3][3];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) Mm[r][c] = J[(3 * i + r) + 12 * (3 * i + c)];
  const double c00 = Mm[1][1] * Mm[2][2] - Mm[1][2] * Mm[2][1];
  const double c10 = Mm[1][2] * Mm[2][0] - Mm[1][0] * Mm[2][2];
  const double c20 = Mm[1][0] * Mm[2][1] - Mm[1][1] * Mm[2][0];
  const double det = c00 * Mm[0][0] + c10 * Mm[0][1] + c20 * Mm[0][2];
  const double id = 1.0 / det;
  I[0][0] = c00 * id;
  I[1][0] = c10 * id;
  I[2][0] = c20 * id;
  I[0][1] = (Mm[0][2] * Mm[2][1] - Mm[0][1] * Mm[2][2]) * id;
  I[1][1] = (Mm[0][0] * Mm[2][2] - Mm[0][2] * Mm[2][0]) * id;
  I[2][1] = (Mm[0][1] * Mm[2][0] - Mm[0][0] * Mm[2][1]) * id;
  I[0][2] = (Mm[0][1] * Mm[1][2] - Mm[0][2] * Mm[1][1]) * id;
  I[1][2] = (Mm[0][2] * Mm[1][0] - Mm[0][0] * Mm[1][2]) * id;
  I[2][2] = (Mm[0][0] * Mm[1][1] - Mm[0][1] * Mm[1][0]) * id;
  for (int r = 0; r < 3; ++r) {
    double s0 = 0, s1 = 0, s2 = 0;
    for (int c = 0; c < 3; ++c) {
      s0 += I[r][c] * a[c];
      s1 += I[r][c] * vg[c];
      s2 += I[r][c] * e[c];
    }
    ddq[o12 + 3 * i + r] = s0;
    dq_cmd[o12 + 3 * i + r] = s1;
    q_step[o12 + 3 * i + r] = s2;
  }
}

// ------------------------------------------------------------------------------------------
// Whole-body tick around the QPWBC solve (scripts/QP_WBC.py:52-131), four threads per robot (one per leg).
//   pre : fixed-base leg kinematics (solo12InvKin.py:47-59) -> InvKin (src/InvKin.cpp:23-73) -> contact Jacobians
//         (QP_WBC.py:96-101) and the base wrench of rnea(q, dq, ddq_cmd) (:104)
//   post: rnea(q, dq, ddq_cmd + [ddq_res; 0])[6:] - Jc[:, 6:]' f_res (:110-117)
// Layouts: q [B][19] (xyz, quaternion xyzw, 12 joints), dq [B][18] (base linear and angular velocity in the base
// frame, 12 joints), contacts [B][4], goals [B][12] (3 x 4 column-major), Jc [B][12*18] column-major.
// ------------------------------------------------------------------------------------------
// One kernel for the whole tick (round 2; round 1 ran pre -> QPWBC -> post as three kernels with Jc / RNEA / ddq handed
// over through HBM: 3.6 KB per robot on top of 1.24 KB algorithmic).  A warp takes GROUPS of G = 2, 4 or 8 consecutive
// robots from a counter (csrc/wbc_tick_core.cuh: the leg code on lanes 4 r + leg, the QP on half-warps with the iterates in
// registers); what the QP needs of the contact Jacobian (its six base columns) and the base wrench stay in the warp's slab.
#ifndef QRW_WT_CTAS
#define QRW_WT_CTAS 4
#endif
constexpr int WT_WARPS = 4, WT_CTAS_PER_SM = QRW_WT_CTAS;
__global__ void __launch_bounds__(32 * WT_WARPS, WT_CTAS_PER_SM)
    wbc_tick_kernel(const WbcConst C, const int B, const int G, const int first, const wt::TickIo io, int* counter) {
  extern __shared__ __align__(16) double slab[];
  double* sl = slab + (threadIdx.x >> 5) * wt::WT2_SLAB_DOUBLES;
  const int lane = threadIdx.x & 31;
  for (;;) {
    int b0 = 0;
    if (lane == 0) b0 = atomicAdd(counter, G);
    b0 = __shfl_sync(0xffffffffu, b0, 0);
    if (b0 >= B) break;
    wt::tick_group(C, B, G, first, io, sl, lane, b0);
  }
}
// pinv(M[:6, :6]) of the batch-wide M, once per handle (same code as the per-robot path: bit-identical)
__global__ void wbc_yinv_kernel(const double* __restrict__ M, double* Yinv) {
  __shared__ double tmp[108];
  if (threadIdx.x == 0) {
    for (int e = 0; e < 36; ++e) tmp[72 + e] = M[(e % 6) + 18 * (e / 6)];
    pinv6_serial(tmp + 72, Yinv, tmp, tmp + 36);
  }
}
// the Pinocchio sub-boundary of scripts/solo12InvKin.py:47-59 on its own (parity tests, drop-in Solo12InvKin):
// posf, vf, wf, af [B][4*3] column-major 4 x 3, Jf [B][12*12] column-major
__global__ void __launch_bounds__(128) solo12_feet_kernel(const int B, const double* __restrict__ q12,
                                                          const double* __restrict__ dq12, double* posf, double* vf,
                                                          double* wf, double* af, double* Jf) {
  using namespace s12;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int leg = t & 3;
  if ((t >> 2) >= B) return;
  const size_t b = (size_t)(t >> 2);
  double qj[3], dqj[3];
  for (int c = 0; c < 3; ++c) {
    qj[c] = q12[b * 12 + 3 * leg + c];
    dqj[c] = dq12[b * 12 + 3 * leg + c];
  }
  const LegKin K = leg_kinematics(leg, qj);
  V3 J[3], v, w, acl;
  leg_jacobian(K, J);
  foot_motion(K, dqj, v, w, acl);
  const V3 asp = acl - cross(w, v);  // spatial acceleration, what getFrameAcceleration returns
  const size_t o = b * 12 + leg;
  posf[o] = K.pf.x, posf[o + 4] = K.pf.y, posf[o + 8] = K.pf.z;
  vf[o] = v.x, vf[o + 4] = v.y, vf[o + 8] = v.z;
  wf[o] = w.x, wf[o + 4] = w.y, wf[o + 8] = w.z;
  af[o] = asp.x, af[o + 4] = asp.y, af[o + 8] = asp.z;
  double* Jb = Jf + b * 144;
  for (int c = 0; c < 12; ++c) {
    const bool mine = c >= 3 * leg && c < 3 * leg + 3;
    const V3 col = mine ? J[c - 3 * leg] : v3(0.0, 0.0, 0.0);
    Jb[3 * leg + 12 * c] = col.x;
    Jb[3 * leg + 1 + 12 * c] = col.y;
    Jb[3 * leg + 2 + 12 * c] = col.z;
  }
}

// FP64 FMA peak: 8 independent DFMA chains per thread, 1024 threads per SM-slot
__global__ void __launch_bounds__(256) fp64_peak_kernel(double* out, int iters, double a, double b) {
  double v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3, v4 = v0 + 4, v5 = v0 + 5, v6 = v0 + 6, v7 = v0 + 7;
  for (int i = 0; i < iters; ++i) {
    v0 = fma(v0, a, b);
    v1 = fma(v1, a, b);
    v2 = fma(v2, a, b);
    v3 = fma(v3, a, b);
    v4 = fma(v4, a, b);
    v5 = fma(v5, a, b);
    v6 = fma(v6, a, b);
    v7 = fma(v7, a, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((v0 + v1) + (v2 + v3)) + ((v4 + v5) + (v6 + v7));
}

// Event behind a launch on the caller's stream, for host getters that copy on the handle's private stream.  Skipped while
// the stream is being captured into a CUDA graph (an event recorded there belongs to the capture): callers that replay
// graphs synchronise the replay stream themselves before reading results.
int record_done(cudaEvent_t ev, cudaStream_t s) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &st) != cudaSuccess) return fail("cudaStreamIsCapturing failed");
  if (st != cudaStreamCaptureStatusNone) return 0;
  CU(cudaEventRecord(ev, s));
  return 0;
}

int sm_count_of(int device) {
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
  return n;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// handles
// ------------------------------------------------------------------------------------------
struct qrw_mpc {
  int B, n_steps, N_gait, device, sms;
  MpcConst C;
  bool dbg_on = false;
  int variant = 2;  // 2 = mpc_v2_kernel (the one shipped); A/B build only: 3 = nested-dissection state solve, 1 = v1 CTA kernel, 0 = warp per instance
  double *d_xref = nullptr, *d_fsteps = nullptr, *d_out = nullptr, *d_state = nullptr, *d_rho = nullptr,
         *d_dbg = nullptr, *d_gait = nullptr, *d_sgait = nullptr;
  int *d_state_i = nullptr, *d_info = nullptr, *d_counter = nullptr, *d_order = nullptr;
  bool use_order = true;       // longest-expected-solve first (mpc_order_kernel); qrw_mpc_set_work_order(h, 0): index order
#ifndef QRW_SPEED_GAIN
#define QRW_SPEED_GAIN 2000.0
#endif
  double speed_gain = QRW_SPEED_GAIN;  // iterations expected per m/s of commanded horizontal speed
  double *h_xref = nullptr, *h_fsteps = nullptr, *h_out = nullptr;  // pinned
  double* h_out_dev = nullptr;   // device-side address of h_out (UVA mapping), nullptr if the mapping is not available
  double* mirror_next = nullptr;  // set by qrw_mpc_run_host around its launch: the kernel mirrors every result into h_out
  bool host_fresh = false;        // h_out holds the results of the latest solve (qrw_mpc_result_host(h, h_out) has nothing to copy)
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;  // recorded behind every solve launch on the CALLER's stream; the host getters wait for it
};
struct qrw_wbc {
  int B, device, sms;
  WbcConst C;
  bool first = true;
  int m_stride = 324;  // 0: one M shared by the whole batch (the fused whole-body tick, where M is a model constant)
  int f_stride = 12;   // doubles between the f_cmd blocks of consecutive robots (24 * n_steps when read from an MPC result)
  double *d_M = nullptr, *d_Jc = nullptr, *d_f = nullptr, *d_R = nullptr, *d_state = nullptr, *d_fres = nullptr,
         *d_ddq = nullptr, *d_H = nullptr;
  int *d_info = nullptr, *d_counter = nullptr;
  cudaStream_t stream = nullptr;
};

// whole-body tick: QPWBC handle + the buffers between its three kernels
struct qrw_wbctick {
  int B, device;
  qrw_wbc* wbc = nullptr;
  double *d_q = nullptr, *d_dq = nullptr, *d_f = nullptr, *d_c = nullptr, *d_pg = nullptr, *d_vg = nullptr,
         *d_ag = nullptr;                                                      // staging of the host entry point
  double *d_Jc = nullptr, *d_rnea = nullptr, *d_ddq = nullptr, *d_vdes = nullptr, *d_qdes = nullptr, *d_tau = nullptr;
  double* d_Yinv = nullptr;  // pinv(M[:6, :6]) of the batch-wide M
  bool debug = false;        // copy Jc / RNEA out of the kernel (parity tests); off: they never leave the SM
  int group = 0;             // robots per warp (2, 4, 8); 0 = chosen from the batch size at launch
};

// diag of crba(q = 0)[:6, :6] (scripts/QP_WBC.py:89-93): total mass, and the moments of inertia of the whole robot
// about the axes of the base frame at the zero configuration
static void solo12_base_inertia(double Y[6]) {
  double m = S12_BASE_MASS, I[3] = {S12_BASE_IXX, S12_BASE_IYY, S12_BASE_IZZ};
  for (int leg = 0; leg < 4; ++leg) {
    const double sx = leg < 2 ? 1.0 : -1.0, sy = (leg & 1) ? -1.0 : 1.0;
    const double o0[3] = {sx * S12_HAA_X, sy * S12_HAA_Y, 0.0};
    const double o1[3] = {o0[0], o0[1] + sy * S12_HFE_Y, o0[2]};
    const double o2[3] = {o1[0], o1[1] + sy * S12_KFE_Y, o1[2] + S12_KFE_Z};
    const double mass[3] = {S12_L0_MASS, S12_L1_MASS, S12_L2_MASS};
    const double com[3][3] = {{o0[0] + sx * S12_L0_CX, o0[1] + sy * S12_L0_CY, o0[2] + S12_L0_CZ},
                              {o1[0] + S12_L1_CX, o1[1] + sy * S12_L1_CY, o1[2] + S12_L1_CZ},
                              {o2[0] + S12_L2_CX, o2[1] + sy * S12_L2_CY, o2[2] + S12_L2_CZ}};
    const double Id[3][3] = {{S12_L0_IXX, S12_L0_IYY, S12_L0_IZZ}, {S12_L1_IXX, S12_L1_IYY, S12_L1_IZZ},
                             {S12_L2_IXX, S12_L2_IYY, S12_L2_IZZ}};
    for (int l = 0; l < 3; ++l) {
      m += mass[l];
      const double r2 = com[l][0] * com[l][0] + com[l][1] * com[l][1] + com[l][2] * com[l][2];
      for (int a = 0; a < 3; ++a) I[a] += Id[l][a] + mass[l] * (r2 - com[l][a] * com[l][a]);
    }
  }
  Y[0] = Y[1] = Y[2] = m;
  Y[3] = I[0], Y[4] = I[1], Y[5] = I[2];
}

extern "C" {

const char* qrw_last_error(void) { return g_err.c_str(); }
long long qrw_launch_count(void) { return g_launches.load(); }
int qrw_device_sm_count(int device, int* sm_count) {
  int n = sm_count_of(device);
  if (n <= 0) return fail("no usable CUDA device (this library has no CPU fallback)");
  *sm_count = n;
  return 0;
}

// Measured DFMA throughput of the device (the FP64 roofline denominator bench.py reports).
int qrw_fp64_peak(int device, double* tflops) {
  if (!tflops) return fail("qrw_fp64_peak: null argument");
  const int sms = sm_count_of(device);
  if (sms <= 0) return fail("qrw_fp64_peak: no usable CUDA device (this library has no CPU fallback)");
  CU(cudaSetDevice(device));
  const int blocks = sms * 8, iters = 1 << 14;
  double* d = nullptr;
  CU(cudaMalloc(&d, (size_t)blocks * 256 * sizeof(double)));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CU(cudaEventRecord(e0));
    fp64_peak_kernel<<<blocks, 256>>>(d, iters, 0.999999, 1e-9);
    ++g_launches;
    CU(cudaEventRecord(e1));
    CU(cudaEventSynchronize(e1));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *tflops = 2.0 * 8.0 * (double)iters * 256.0 * blocks / (best * 1e-3) / 1e12;
  return 0;
}

// ---------------------------------------------------------------- MPC
int qrw_mpc_create(qrw_mpc_t** out, int batch, double dt, int n_steps, double T_gait, int N_gait, int device) {
  (void)T_gait;  // stored by the reference (MPC.cpp:6) but never read on this path
  if (!out) return fail("qrw_mpc_create: null output handle");
  if (batch <= 0) return fail("qrw_mpc_create: batch must be positive");
  if (n_steps != NS)
    return fail("qrw_mpc_create: n_steps = " + std::to_string(n_steps) +
                " is not supported: this library compiles the reference configuration n_steps == 16 only "
                "(T_mpc / dt_mpc = 0.32 / 0.02, src/config_solo12.yaml:11-13; one horizon stage per 16 lanes of a 256-thread CTA)");
  if (N_gait <= n_steps || N_gait > NGMAX)
    return fail("qrw_mpc_create: need n_steps < N_gait <= 32 (the reference writes gait.row(n_active))");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail("qrw_mpc_create: no CUDA device (this library has no CPU fallback)");
  if (device < 0 || device >= ndev) return fail("qrw_mpc_create: bad device index");
  CU(cudaSetDevice(device));
  auto* h = new qrw_mpc();
  h->B = batch;
  h->n_steps = n_steps;
  h->N_gait = N_gait;
  h->device = device;
  h->sms = sm_count_of(device);
  h->C = make_mpc_const(dt, N_gait);
  const size_t B = (size_t)batch;
  CUH(cudaMalloc(&h->d_xref, B * 12 * (NS + 1) * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_fsteps, B * 12 * N_gait * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_out, B * 24 * NS * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_state, B * MPC_STATE_DOUBLES * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_state_i, B * MPC_STATE_INTS * sizeof(int)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_rho, B * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_info, B * 4 * sizeof(int)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_counter, sizeof(int)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_order, B * sizeof(int)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_gait, B * 4 * N_gait * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMalloc(&h->d_sgait, B * 12 * NS * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMemset(h->d_state, 0, B * MPC_STATE_DOUBLES * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMemset(h->d_state_i, 0, B * MPC_STATE_INTS * sizeof(int)), qrw_mpc_destroy(h));
  CUH(cudaMemset(h->d_out, 0, B * 24 * NS * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMemset(h->d_info, 0, B * 4 * sizeof(int)), qrw_mpc_destroy(h));
  CUH(cudaMemset(h->d_rho, 0, B * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMallocHost(&h->h_xref, B * 12 * (NS + 1) * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMallocHost(&h->h_fsteps, B * 12 * N_gait * sizeof(double)), qrw_mpc_destroy(h));
  CUH(cudaMallocHost(&h->h_out, B * 24 * NS * sizeof(double)), qrw_mpc_destroy(h));
  {  // under unified addressing pinned host memory is mapped into the device's address space: kernels can write it
    void* dp = nullptr;
    if (cudaHostGetDevicePointer(&dp, h->h_out, 0) == cudaSuccess) h->h_out_dev = static_cast<double*>(dp);
    else cudaGetLastError();
  }
  CUH(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), qrw_mpc_destroy(h));
  CUH(cudaEventCreateWithFlags(&h->done, cudaEventDisableTiming), qrw_mpc_destroy(h));
  CUH(cudaFuncSetAttribute(mpc_v2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, v2::V2_SLAB_BYTES), qrw_mpc_destroy(h));
#ifdef QRW_AB_KERNELS
  CUH(cudaFuncSetAttribute(mpc_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MPC_SLAB_BYTES), qrw_mpc_destroy(h));
  CUH(cudaFuncSetAttribute(mpc_cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, cta::CTA_SLAB_BYTES), qrw_mpc_destroy(h));
  CUH(cudaFuncSetAttribute(mpc_v3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, nd::V3_SLAB_BYTES), qrw_mpc_destroy(h));
  CUH(cudaFuncSetAttribute(mpc_v4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, v4::V4_SLAB_BYTES), qrw_mpc_destroy(h));
#endif
  *out = h;
  return 0;
}

int qrw_mpc_destroy(qrw_mpc_t* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaFree(h->d_xref);
  cudaFree(h->d_fsteps);
  cudaFree(h->d_out);
  cudaFree(h->d_state);
  cudaFree(h->d_state_i);
  cudaFree(h->d_rho);
  cudaFree(h->d_info);
  cudaFree(h->d_counter);
  cudaFree(h->d_order);
  cudaFree(h->d_gait);
  cudaFree(h->d_sgait);
  cudaFree(h->d_dbg);
  cudaFreeHost(h->h_xref);
  cudaFreeHost(h->h_fsteps);
  cudaFreeHost(h->h_out);
  if (h->done) cudaEventDestroy(h->done);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

int qrw_mpc_run_device(qrw_mpc_t* h, int num_iter, const double* d_xref, const double* d_fsteps, void* stream) {
  if (!h || !d_xref || !d_fsteps) return fail("qrw_mpc_run_device: null argument");
  CU(cudaSetDevice(h->device));
  h->host_fresh = false;
  cudaStream_t s = (cudaStream_t)stream;
  CU(cudaMemsetAsync(h->d_counter, 0, sizeof(int), s));
#ifdef QRW_AB_KERNELS
  if (h->variant == 4) {
    int grid = 2 * h->sms;
    if (grid > h->B) grid = h->B;
    mpc_v4_kernel<<<grid, v4::NT, v4::V4_SLAB_BYTES, s>>>(h->C, h->B, num_iter == 0 ? 1 : 0, h->sms, d_xref, d_fsteps,
                                                          h->d_state, h->d_state_i, h->d_out, h->d_info, h->d_rho,
                                                          h->dbg_on ? h->d_dbg : nullptr, h->d_counter, nullptr, nullptr);
  } else if (h->variant == 3) {
    int grid = 2 * h->sms;
    if (grid > h->B) grid = h->B;
    mpc_v3_kernel<<<grid, nd::NT, nd::V3_SLAB_BYTES, s>>>(h->C, h->B, num_iter == 0 ? 1 : 0, d_xref, d_fsteps, h->d_state,
                                                          h->d_state_i, h->d_out, h->d_info, h->d_rho,
                                                          h->dbg_on ? h->d_dbg : nullptr, h->d_counter, nullptr, nullptr);
  } else if (h->variant == 1) {
    int grid = 2 * h->sms;  // two 84 KB slabs (16 warps) per SM
    if (grid > h->B) grid = h->B;
    mpc_cta_kernel<<<grid, cta::NT, cta::CTA_SLAB_BYTES, s>>>(h->C, h->B, num_iter == 0 ? 1 : 0, d_xref, d_fsteps,
                                                              h->d_state, h->d_state_i, h->d_out, h->d_info,
                                                              h->d_rho, h->dbg_on ? h->d_dbg : nullptr,
                                                              h->d_counter);
  } else if (h->variant == 0) {
    int grid = 3 * h->sms;  // three 75 KB slabs per SM
    if (grid > h->B) grid = h->B;
    mpc_solve_kernel<<<grid, 32, MPC_SLAB_BYTES, s>>>(h->C, h->B, num_iter == 0 ? 1 : 0, d_xref, d_fsteps,
                                                      h->d_state, h->d_state_i, h->d_out, h->d_info, h->d_rho,
                                                      h->dbg_on ? h->d_dbg : nullptr, h->d_counter);
  } else
#endif
  {
    int grid = 2 * h->sms;  // two CTAs (16 warps, 128 registers per thread) per SM
    if (grid > h->B) grid = h->B;
    const bool ord = h->use_order && h->B > grid;  // the order only matters when CTAs take more than one robot
    if (ord) {
      mpc_order_kernel<<<1, 1024, 0, s>>>(h->d_info, d_xref, h->speed_gain, h->B, h->d_order);
      ++g_launches;
    }
    mpc_v2_kernel<<<grid, v2::NT, v2::V2_SLAB_BYTES, s>>>(h->C, h->B, num_iter == 0 ? 1 : 0, h->sms, d_xref, d_fsteps,
                                                          h->d_state, h->d_state_i, h->d_out, h->d_info, h->d_rho,
                                                          h->dbg_on ? h->d_dbg : nullptr, h->d_counter, nullptr,
                                                          ord ? h->d_order : nullptr, h->mirror_next);
  }
  ++g_launches;
  CU(cudaGetLastError());
  return record_done(h->done, s);  // qrw_mpc_result_host / gait_host / sgait_host order themselves behind this launch
}

// MPC::run with a PER-ROBOT call index: d_k[b] < 0 -> robot b is not called in this launch (its solver state and its
// result stay as they are), 0 -> cold start (create_matrices / osqp_setup path), > 0 -> update path.  This is what a
// server needs when robots post their MPC requests at different times (qrw_b200/mpc_server.py).
int qrw_mpc_run_indexed_device(qrw_mpc_t* h, const int* d_k, const double* d_xref, const double* d_fsteps, void* stream) {
  if (!h || !d_k || !d_xref || !d_fsteps) return fail("qrw_mpc_run_indexed_device: null argument");
  if (h->variant != 2) return fail("qrw_mpc_run_indexed_device: only the default kernel (variant 2) takes a per-robot index");
  CU(cudaSetDevice(h->device));
  h->host_fresh = false;
  cudaStream_t s = (cudaStream_t)stream;
  CU(cudaMemsetAsync(h->d_counter, 0, sizeof(int), s));
  int grid = 2 * h->sms;
  if (grid > h->B) grid = h->B;
  mpc_v2_kernel<<<grid, v2::NT, v2::V2_SLAB_BYTES, s>>>(h->C, h->B, 0, h->sms, d_xref, d_fsteps, h->d_state, h->d_state_i, h->d_out,
                                                        h->d_info, h->d_rho, h->dbg_on ? h->d_dbg : nullptr, h->d_counter, d_k,
                                                        nullptr, nullptr);
  ++g_launches;
  CU(cudaGetLastError());
  return record_done(h->done, s);
}

// 1 (default) = longest-expected-solve first (mpc_order_kernel), 0 = robots are taken in index order
int qrw_mpc_set_work_order(qrw_mpc_t* h, int longest_first) {
  if (!h) return fail("qrw_mpc_set_work_order: null handle");
  h->use_order = longest_first != 0;
  return 0;
}

// device-side address of a pinned (page-locked, UVA-mapped) host pointer; nullptr for pageable memory
static const double* mapped_host_ptr(const double* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return (a.type == cudaMemoryTypeHost && a.devicePointer) ? static_cast<const double*>(a.devicePointer) : nullptr;
}

// Host entry point.  Pageable inputs are staged with two copies.  PINNED inputs (cudaMallocHost / cudaHostRegister / torch
// pin_memory: e.g. the handle's own buffers, qrw_mpc_pinned) are not copied at all: the kernel reads each robot's xref / fsteps
// straight from host memory when it starts on that robot (3.5 KB over PCIe against a ~6 ms solve) and mirrors the robot's result
// into the handle's pinned result buffer when it is done, so both transfers run under the solves of the other robots instead of
// in front of and behind the launch.
int qrw_mpc_run_host(qrw_mpc_t* h, int num_iter, const double* xref, const double* fsteps) {
  if (!h || !xref || !fsteps) return fail("qrw_mpc_run_host: null argument");
  CU(cudaSetDevice(h->device));
  const size_t B = (size_t)h->B;
  const double* mx = h->variant == 2 ? mapped_host_ptr(xref) : nullptr;
  const double* mf = mx ? mapped_host_ptr(fsteps) : nullptr;
  const bool zero_copy = mx && mf;
  if (!zero_copy) {
    CU(cudaMemcpyAsync(h->d_xref, xref, B * 12 * (NS + 1) * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(h->d_fsteps, fsteps, B * 12 * h->N_gait * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  }
  h->mirror_next = zero_copy ? h->h_out_dev : nullptr;
  const bool mirrored = h->mirror_next != nullptr;
  const int rc = qrw_mpc_run_device(h, num_iter, zero_copy ? mx : h->d_xref, zero_copy ? mf : h->d_fsteps, h->stream);
  h->mirror_next = nullptr;
  if (rc) return 1;
  CU(cudaStreamSynchronize(h->stream));
  h->host_fresh = mirrored;
  return 0;
}

int qrw_mpc_result_host(qrw_mpc_t* h, double* out) {
  if (!h || !out) return fail("qrw_mpc_result_host: null argument");
  CU(cudaSetDevice(h->device));
  if (h->host_fresh && out == h->h_out) return 0;  // the kernel has written it there already (qrw_mpc_run_host, pinned inputs)
  CU(cudaStreamWaitEvent(h->stream, h->done, 0));  // the last solve may have been launched on another stream
  CU(cudaMemcpyAsync(out, h->d_out, (size_t)h->B * 24 * NS * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int qrw_mpc_result_device(qrw_mpc_t* h, const double** d_out) {
  if (!h || !d_out) return fail("qrw_mpc_result_device: null argument");
  *d_out = h->d_out;
  return 0;
}

static int mpc_unpack(qrw_mpc_t* h) {
  CU(cudaStreamWaitEvent(h->stream, h->done, 0));
  const int per = 4 * h->N_gait + 12 * NS;
  const long long total = (long long)h->B * per;
  mpc_unpack_gait_kernel<<<(unsigned)((total + 255) / 256), 256, 0, h->stream>>>(h->d_state_i, h->B, h->N_gait,
                                                                                h->d_gait, h->d_sgait);
  ++g_launches;
  CU(cudaGetLastError());
  return 0;
}
int qrw_mpc_gait_host(qrw_mpc_t* h, double* out) {
  if (!h || !out) return fail("qrw_mpc_gait_host: null argument");
  CU(cudaSetDevice(h->device));
  if (mpc_unpack(h)) return 1;
  CU(cudaMemcpyAsync(out, h->d_gait, (size_t)h->B * 4 * h->N_gait * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int qrw_mpc_sgait_host(qrw_mpc_t* h, double* out) {
  if (!h || !out) return fail("qrw_mpc_sgait_host: null argument");
  CU(cudaSetDevice(h->device));
  if (mpc_unpack(h)) return 1;
  CU(cudaMemcpyAsync(out, h->d_sgait, (size_t)h->B * 12 * NS * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}

int qrw_mpc_info_host(qrw_mpc_t* h, int* iters, int* status, double* rho, int* n_factor) {
  if (!h) return fail("qrw_mpc_info_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  int* tmp = new int[(size_t)h->B * 4];
  cudaError_t e = cudaMemcpy(tmp, h->d_info, (size_t)h->B * 4 * sizeof(int), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && rho) e = cudaMemcpy(rho, h->d_rho, (size_t)h->B * sizeof(double), cudaMemcpyDeviceToHost);
  for (int b = 0; b < h->B; ++b) {
    if (iters) iters[b] = tmp[4 * b];
    if (status) status[b] = tmp[4 * b + 1];
    if (n_factor) n_factor[b] = tmp[4 * b + 2];
  }
  delete[] tmp;
  if (e != cudaSuccess) return fail(std::string("qrw_mpc_info_host: ") + cudaGetErrorString(e));
  return 0;
}

int qrw_mpc_set_max_iter(qrw_mpc_t* h, int max_iter) {
  if (!h) return fail("qrw_mpc_set_max_iter: null handle");
  if (max_iter <= 0 || max_iter % h->C.check_every != 0)
    return fail("qrw_mpc_set_max_iter: max_iter must be a positive multiple of 25 (check_termination)");
  h->C.max_iter = max_iter;
  return 0;
}
int qrw_mpc_set_kernel(qrw_mpc_t* h, int variant) {
  if (!h) return fail("qrw_mpc_set_kernel: null handle");
#ifdef QRW_AB_KERNELS
  if (variant < 0 || variant > 4)
    return fail("qrw_mpc_set_kernel: variant must be 4 (super-stage chain), 0 (warp per instance), 1 (CTA per instance, v1), 2 (force-first CTA, twisted state chain: default) or 3 (nested-dissection state solve)");
#else
  if (variant != 2)
    return fail("qrw_mpc_set_kernel: this library contains the one MPC kernel it ships (variant 2); variants 0, 1 and 3 exist only in the A/B "
                "test build (-DQRW_AB_KERNELS, tests/ab/libqrw_b200_ab.so)");
#endif
  h->variant = variant;
  return 0;
}
int qrw_mpc_debug_host(qrw_mpc_t* h, double* out) {
  if (!h) return fail("qrw_mpc_debug_host: null handle");
  CU(cudaSetDevice(h->device));
  if (!h->d_dbg) {
    CU(cudaMalloc(&h->d_dbg, (size_t)h->B * 1281 * sizeof(double)));
    CU(cudaMemset(h->d_dbg, 0, (size_t)h->B * 1281 * sizeof(double)));
  }
  h->dbg_on = true;
  if (out) {
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(out, h->d_dbg, (size_t)h->B * 1281 * sizeof(double), cudaMemcpyDeviceToHost));
  }
  return 0;
}

int qrw_mpc_state_bytes(qrw_mpc_t* h, size_t* bytes) {
  if (!h || !bytes) return fail("qrw_mpc_state_bytes: null argument");
  *bytes = (size_t)h->B * (MPC_STATE_DOUBLES * sizeof(double) + MPC_STATE_INTS * sizeof(int));
  return 0;
}
int qrw_mpc_state_save(qrw_mpc_t* h, void* d_buf, void* stream) {
  if (!h || !d_buf) return fail("qrw_mpc_state_save: null argument");
  CU(cudaSetDevice(h->device));
  const size_t nd = (size_t)h->B * MPC_STATE_DOUBLES * sizeof(double), ni = (size_t)h->B * MPC_STATE_INTS * sizeof(int);
  CU(cudaMemcpyAsync(d_buf, h->d_state, nd, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  CU(cudaMemcpyAsync((char*)d_buf + nd, h->d_state_i, ni, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}
int qrw_mpc_state_load(qrw_mpc_t* h, const void* d_buf, void* stream) {
  if (!h || !d_buf) return fail("qrw_mpc_state_load: null argument");
  CU(cudaSetDevice(h->device));
  const size_t nd = (size_t)h->B * MPC_STATE_DOUBLES * sizeof(double), ni = (size_t)h->B * MPC_STATE_INTS * sizeof(int);
  CU(cudaMemcpyAsync(h->d_state, d_buf, nd, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  CU(cudaMemcpyAsync(h->d_state_i, (const char*)d_buf + nd, ni, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}
int qrw_mpc_pinned(qrw_mpc_t* h, double** xref, double** fsteps, double** result) {
  if (!h) return fail("qrw_mpc_pinned: null handle");
  if (xref) *xref = h->h_xref;
  if (fsteps) *fsteps = h->h_fsteps;
  if (result) *result = h->h_out;
  return 0;
}

// ---------------------------------------------------------------- QPWBC
int qrw_wbc_create(qrw_wbc_t** out, int batch, int device) {
  if (!out) return fail("qrw_wbc_create: null output handle");
  if (batch <= 0) return fail("qrw_wbc_create: batch must be positive");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail("qrw_wbc_create: no CUDA device (this library has no CPU fallback)");
  if (device < 0 || device >= ndev) return fail("qrw_wbc_create: bad device index");
  CU(cudaSetDevice(device));
  auto* h = new qrw_wbc();
  h->B = batch;
  h->device = device;
  h->sms = sm_count_of(device);
  h->C = make_wbc_const();
  const size_t B = (size_t)batch;
  CUH(cudaMalloc(&h->d_M, B * 324 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_Jc, B * 216 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_f, B * 12 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_R, B * 6 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_state, B * WBC_STATE_DOUBLES * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_fres, B * 12 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_ddq, B * 6 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_H, B * 144 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_info, B * 4 * sizeof(int)), qrw_wbc_destroy(h));
  CUH(cudaMalloc(&h->d_counter, sizeof(int)), qrw_wbc_destroy(h));
  CUH(cudaMemset(h->d_state, 0, B * WBC_STATE_DOUBLES * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMemset(h->d_fres, 0, B * 12 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMemset(h->d_ddq, 0, B * 6 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMemset(h->d_H, 0, B * 144 * sizeof(double)), qrw_wbc_destroy(h));
  CUH(cudaMemset(h->d_info, 0, B * 4 * sizeof(int)), qrw_wbc_destroy(h));
  CUH(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), qrw_wbc_destroy(h));
  *out = h;
  return 0;
}
int qrw_wbc_destroy(qrw_wbc_t* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaFree(h->d_M);
  cudaFree(h->d_Jc);
  cudaFree(h->d_f);
  cudaFree(h->d_R);
  cudaFree(h->d_state);
  cudaFree(h->d_fres);
  cudaFree(h->d_ddq);
  cudaFree(h->d_H);
  cudaFree(h->d_info);
  cudaFree(h->d_counter);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}
int qrw_wbc_run_device(qrw_wbc_t* h, const double* d_M, const double* d_Jc, const double* d_f_cmd,
                       const double* d_RNEA, void* stream) {
  if (!h || !d_M || !d_Jc || !d_f_cmd || !d_RNEA) return fail("qrw_wbc_run_device: null argument");
  CU(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  CU(cudaMemsetAsync(h->d_counter, 0, sizeof(int), s));
  int grid = 7 * h->sms;  // 7 CTAs x 4 warps x 7.9 KB per SM
  const int need = (h->B + WBC_WARPS - 1) / WBC_WARPS;
  if (grid > need) grid = need;
  wbc_solve_kernel<<<grid, 32 * WBC_WARPS, WBC_WARPS * WBC_SLAB_BYTES, s>>>(
      h->C, h->B, h->first ? 1 : 0, d_M, d_Jc, d_f_cmd, d_RNEA, h->d_state, h->d_fres, h->d_ddq, h->d_H, h->d_info,
      h->d_counter, h->m_stride, h->f_stride);
  ++g_launches;
  CU(cudaGetLastError());
  h->first = false;  // QPWBC::initialized, src/QPWBC.cpp:23,254
  return 0;
}
int qrw_wbc_run_host(qrw_wbc_t* h, const double* M, const double* Jc, const double* f_cmd, const double* RNEA) {
  if (!h || !M || !Jc || !f_cmd || !RNEA) return fail("qrw_wbc_run_host: null argument");
  CU(cudaSetDevice(h->device));
  const size_t B = (size_t)h->B;
  CU(cudaMemcpyAsync(h->d_M, M, B * 324 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  CU(cudaMemcpyAsync(h->d_Jc, Jc, B * 216 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  CU(cudaMemcpyAsync(h->d_f, f_cmd, B * 12 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  CU(cudaMemcpyAsync(h->d_R, RNEA, B * 6 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  if (qrw_wbc_run_device(h, h->d_M, h->d_Jc, h->d_f, h->d_R, h->stream)) return 1;
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int qrw_wbc_result_host(qrw_wbc_t* h, double* f_res, double* ddq_res, double* H) {
  if (!h) return fail("qrw_wbc_result_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  const size_t B = (size_t)h->B;
  if (f_res) CU(cudaMemcpy(f_res, h->d_fres, B * 12 * sizeof(double), cudaMemcpyDeviceToHost));
  if (ddq_res) CU(cudaMemcpy(ddq_res, h->d_ddq, B * 6 * sizeof(double), cudaMemcpyDeviceToHost));
  if (H) CU(cudaMemcpy(H, h->d_H, B * 144 * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}
int qrw_wbc_result_device(qrw_wbc_t* h, const double** d_f_res, const double** d_ddq_res) {
  if (!h) return fail("qrw_wbc_result_device: null handle");
  if (d_f_res) *d_f_res = h->d_fres;
  if (d_ddq_res) *d_ddq_res = h->d_ddq;
  return 0;
}
int qrw_wbc_info_host(qrw_wbc_t* h, int* iters, int* status, double* rho) {
  if (!h) return fail("qrw_wbc_info_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  const size_t B = (size_t)h->B;
  int* tmp = new int[B * 4];
  double* st = rho ? new double[B * WBC_STATE_DOUBLES] : nullptr;
  cudaError_t e = cudaMemcpy(tmp, h->d_info, B * 4 * sizeof(int), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && rho) e = cudaMemcpy(st, h->d_state, B * WBC_STATE_DOUBLES * sizeof(double), cudaMemcpyDeviceToHost);
  for (size_t b = 0; b < B; ++b) {
    if (iters) iters[b] = tmp[4 * b];
    if (status) status[b] = tmp[4 * b + 1];
    if (rho) rho[b] = st[b * WBC_STATE_DOUBLES + WS_RHO];
  }
  delete[] tmp;
  delete[] st;
  if (e != cudaSuccess) return fail(std::string("qrw_wbc_info_host: ") + cudaGetErrorString(e));
  return 0;
}

// ---------------------------------------------------------------- InvKin
int qrw_invkin_run_device(int batch, const double* contacts, const double* goals, const double* vgoals,
                          const double* agoals, const double* posf, const double* vf, const double* wf,
                          const double* af, const double* Jf, double* ddq, double* dq_cmd, double* q_step,
                          void* stream) {
  if (batch <= 0) return fail("qrw_invkin_run_device: batch must be positive");
  if (!contacts || !goals || !vgoals || !agoals || !posf || !vf || !wf || !af || !Jf || !ddq || !dq_cmd || !q_step)
    return fail("qrw_invkin_run_device: null argument");
  const long long threads = 4LL * batch;
  invkin_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      batch, contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf, ddq, dq_cmd, q_step);
  ++g_launches;
  CU(cudaGetLastError());
  return 0;
}
int qrw_invkin_run_host(int batch, const double* contacts, const double* goals, const double* vgoals,
                        const double* agoals, const double* posf, const double* vf, const double* wf,
                        const double* af, const double* Jf, double* ddq, double* dq_cmd, double* q_step) {
  if (batch <= 0) return fail("qrw_invkin_run_host: batch must be positive");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail("qrw_invkin_run_host: no CUDA device (this library has no CPU fallback)");
  const size_t B = (size_t)batch;
  const size_t in_sz[9] = {4, 12, 12, 12, 12, 12, 12, 12, 144};
  const double* in_h[9] = {contacts, goals, vgoals, agoals, posf, vf, wf, af, Jf};
  double* in_d[9] = {nullptr};
  double* out_d[3] = {nullptr};
  double* out_h[3] = {ddq, dq_cmd, q_step};
  int rc = 0;
  for (int i = 0; i < 9 && !rc; ++i) {
    if (cudaMalloc(&in_d[i], B * in_sz[i] * sizeof(double)) != cudaSuccess ||
        cudaMemcpy(in_d[i], in_h[i], B * in_sz[i] * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess)
      rc = fail("qrw_invkin_run_host: allocation / H2D copy failed");
  }
  for (int i = 0; i < 3 && !rc; ++i)
    if (cudaMalloc(&out_d[i], B * 12 * sizeof(double)) != cudaSuccess) rc = fail("qrw_invkin_run_host: allocation failed");
  if (!rc)
    rc = qrw_invkin_run_device(batch, in_d[0], in_d[1], in_d[2], in_d[3], in_d[4], in_d[5], in_d[6], in_d[7], in_d[8],
                               out_d[0], out_d[1], out_d[2], nullptr);
  for (int i = 0; i < 3 && !rc; ++i)
    if (cudaMemcpy(out_h[i], out_d[i], B * 12 * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess)
      rc = fail("qrw_invkin_run_host: D2H copy failed");
  for (int i = 0; i < 9; ++i) cudaFree(in_d[i]);
  for (int i = 0; i < 3; ++i) cudaFree(out_d[i]);
  return rc;
}

// ---------------------------------------------------------------- whole-body tick (QP_WBC.py:52-131)
int qrw_wbctick_create(qrw_wbctick_t** out, int batch, int device) {
  if (!out) return fail("qrw_wbctick_create: null output handle");
  qrw_wbc_t* w = nullptr;
  if (qrw_wbc_create(&w, batch, device)) return 1;
  auto* h = new qrw_wbctick();
  h->B = batch;
  h->device = device;
  h->wbc = w;
  const size_t B = (size_t)batch;
  w->m_stride = 0;  // M is a model constant here: one copy for the whole batch
  double M[324] = {0.0}, Y[6];
  solo12_base_inertia(Y);
  for (int i = 0; i < 6; ++i) M[i + 18 * i] = Y[i];
  CUH(cudaMemcpy(w->d_M, M, sizeof(M), cudaMemcpyHostToDevice), qrw_wbctick_destroy(h));
  const size_t sz[13] = {19, 18, 12, 4, 12, 12, 12, 216, 6, 12, 12, 12, 12};
  double** p[13] = {&h->d_q, &h->d_dq, &h->d_f, &h->d_c, &h->d_pg, &h->d_vg, &h->d_ag, &h->d_Jc, &h->d_rnea, &h->d_ddq,
                    &h->d_vdes, &h->d_qdes, &h->d_tau};
  for (int i = 0; i < 13; ++i) {
    CUH(cudaMalloc(p[i], B * sz[i] * sizeof(double)), qrw_wbctick_destroy(h));
    CUH(cudaMemset(*p[i], 0, B * sz[i] * sizeof(double)), qrw_wbctick_destroy(h));
  }
  CUH(cudaMalloc(&h->d_Yinv, 36 * sizeof(double)), qrw_wbctick_destroy(h));
  wbc_yinv_kernel<<<1, 32>>>(w->d_M, h->d_Yinv);
  ++g_launches;
  CUH(cudaGetLastError(), qrw_wbctick_destroy(h));
  CUH(cudaFuncSetAttribute(wbc_tick_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WT_WARPS * wt::WT2_SLAB_BYTES), qrw_wbctick_destroy(h));
  CUH(cudaDeviceSynchronize(), qrw_wbctick_destroy(h));
  *out = h;
  return 0;
}
// 1: every tick also copies the intermediates handed to QPWBC (Jc, RNEA) to HBM for qrw_wbctick_debug_host; 0 (default): off
int qrw_wbctick_set_debug(qrw_wbctick_t* h, int on) {
  if (!h) return fail("qrw_wbctick_set_debug: null handle");
  h->debug = on != 0;
  return 0;
}
// Robots per warp of the tick kernel: 2, 4 or 8; 0 (default) = chosen at launch from the batch size (8 once that fills every
// warp the GPU can hold, smaller groups for small batches).  Results do not depend on it.
int qrw_wbctick_set_group(qrw_wbctick_t* h, int group) {
  if (!h) return fail("qrw_wbctick_set_group: null handle");
  if (group != 0 && group != 2 && group != 4 && group != 8) return fail("qrw_wbctick_set_group: group must be 0, 2, 4 or 8");
  h->group = group;
  return 0;
}
int qrw_wbctick_destroy(qrw_wbctick_t* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  double* p[13] = {h->d_q, h->d_dq, h->d_f, h->d_c, h->d_pg, h->d_vg, h->d_ag, h->d_Jc, h->d_rnea, h->d_ddq, h->d_vdes,
                   h->d_qdes, h->d_tau};
  for (double* x : p) cudaFree(x);
  cudaFree(h->d_Yinv);
  qrw_wbc_destroy(h->wbc);
  delete h;
  return 0;
}
int qrw_wbctick_run_device(qrw_wbctick_t* h, const double* q, const double* dq, const double* f_cmd,
                           const double* contacts, const double* pgoals, const double* vgoals, const double* agoals,
                           void* stream) {
  if (!h || !q || !dq || !f_cmd || !contacts || !pgoals || !vgoals || !agoals)
    return fail("qrw_wbctick_run_device: null argument");
  CU(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  qrw_wbc* w = h->wbc;
  CU(cudaMemsetAsync(w->d_counter, 0, sizeof(int), s));
  // group size: eight robots per warp keep every lane of the leg code busy; a batch too small to fill the GPU's warps that way
  // takes smaller groups (more warps in flight, a shorter critical path per warp)
  const int slots = WT_CTAS_PER_SM * w->sms * WT_WARPS;
  const int G = h->group ? h->group : (h->B >= 8 * slots ? 8 : (h->B >= 4 * slots ? 4 : 2));
  int grid = WT_CTAS_PER_SM * w->sms;  // persistent
  const int need = ((h->B + G - 1) / G + WT_WARPS - 1) / WT_WARPS;
  if (grid > need) grid = need;
  wt::TickIo io{q, dq, f_cmd, contacts, pgoals, vgoals, agoals, h->d_Yinv, w->f_stride, w->d_state, w->d_fres, w->d_ddq, w->d_info,
                h->d_ddq, h->d_vdes, h->d_qdes, h->d_tau, h->debug ? h->d_Jc : nullptr, h->debug ? h->d_rnea : nullptr};
  wbc_tick_kernel<<<grid, 32 * WT_WARPS, WT_WARPS * wt::WT2_SLAB_BYTES, s>>>(w->C, h->B, G, w->first ? 1 : 0, io, w->d_counter);
  ++g_launches;
  CU(cudaGetLastError());
  w->first = false;  // QPWBC::initialized, src/QPWBC.cpp:23,254
  return 0;
}
// Same tick with f_cmd read with a stride: f_cmd + b * f_stride (f_stride >= 12), e.g. the first-step contact forces
// x_f_mpc[12:24, 0] straight out of an MPC result: qrw_mpc_result_device + 12 with stride 24 * n_steps.
int qrw_wbctick_run_fstride_device(qrw_wbctick_t* h, const double* q, const double* dq, const double* f_cmd, int f_stride,
                                   const double* contacts, const double* pgoals, const double* vgoals, const double* agoals,
                                   void* stream) {
  if (!h) return fail("qrw_wbctick_run_fstride_device: null handle");
  if (f_stride < 12) return fail("qrw_wbctick_run_fstride_device: f_stride < 12");
  h->wbc->f_stride = f_stride;
  const int rc = qrw_wbctick_run_device(h, q, dq, f_cmd, contacts, pgoals, vgoals, agoals, stream);
  h->wbc->f_stride = 12;
  return rc;
}
int qrw_wbctick_run_host(qrw_wbctick_t* h, const double* q, const double* dq, const double* f_cmd, const double* contacts,
                         const double* pgoals, const double* vgoals, const double* agoals) {
  if (!h || !q || !dq || !f_cmd || !contacts || !pgoals || !vgoals || !agoals)
    return fail("qrw_wbctick_run_host: null argument");
  CU(cudaSetDevice(h->device));
  const size_t B = (size_t)h->B;
  cudaStream_t s = h->wbc->stream;
  const size_t sz[7] = {19, 18, 12, 4, 12, 12, 12};
  const double* src[7] = {q, dq, f_cmd, contacts, pgoals, vgoals, agoals};
  double* dst[7] = {h->d_q, h->d_dq, h->d_f, h->d_c, h->d_pg, h->d_vg, h->d_ag};
  for (int i = 0; i < 7; ++i) CU(cudaMemcpyAsync(dst[i], src[i], B * sz[i] * sizeof(double), cudaMemcpyHostToDevice, s));
  if (qrw_wbctick_run_device(h, h->d_q, h->d_dq, h->d_f, h->d_c, h->d_pg, h->d_vg, h->d_ag, s)) return 1;
  CU(cudaStreamSynchronize(s));
  return 0;
}
int qrw_wbctick_result_host(qrw_wbctick_t* h, double* tau_ff, double* qdes, double* vdes, double* f_res, double* ddq_cmd) {
  if (!h) return fail("qrw_wbctick_result_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  const size_t n = (size_t)h->B * 12 * sizeof(double);
  if (tau_ff) CU(cudaMemcpy(tau_ff, h->d_tau, n, cudaMemcpyDeviceToHost));
  if (qdes) CU(cudaMemcpy(qdes, h->d_qdes, n, cudaMemcpyDeviceToHost));
  if (vdes) CU(cudaMemcpy(vdes, h->d_vdes, n, cudaMemcpyDeviceToHost));
  if (f_res) CU(cudaMemcpy(f_res, h->wbc->d_fres, n, cudaMemcpyDeviceToHost));
  if (ddq_cmd) CU(cudaMemcpy(ddq_cmd, h->d_ddq, n, cudaMemcpyDeviceToHost));
  return 0;
}
int qrw_wbctick_result_device(qrw_wbctick_t* h, const double** tau_ff, const double** qdes, const double** vdes,
                              const double** f_res) {
  if (!h) return fail("qrw_wbctick_result_device: null handle");
  if (tau_ff) *tau_ff = h->d_tau;
  if (qdes) *qdes = h->d_qdes;
  if (vdes) *vdes = h->d_vdes;
  if (f_res) *f_res = h->wbc->d_fres;
  return 0;
}
int qrw_wbctick_debug_host(qrw_wbctick_t* h, double* Jc, double* rnea6, double* ddq_res) {
  if (!h) return fail("qrw_wbctick_debug_host: null handle");
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  const size_t B = (size_t)h->B;
  if (Jc) CU(cudaMemcpy(Jc, h->d_Jc, B * 216 * sizeof(double), cudaMemcpyDeviceToHost));
  if (rnea6) CU(cudaMemcpy(rnea6, h->d_rnea, B * 6 * sizeof(double), cudaMemcpyDeviceToHost));
  if (ddq_res) CU(cudaMemcpy(ddq_res, h->wbc->d_ddq, B * 6 * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}
int qrw_solo12_base_inertia(double* Y6) {
  if (!Y6) return fail("qrw_solo12_base_inertia: null argument");
  solo12_base_inertia(Y6);
  return 0;
}
// Fixed-base foot kinematics (scripts/solo12InvKin.py:47-59): q12, dq12 [B][12] -> posf, vf, wf, af [B][12]
// (4 x 3 column-major), Jf [B][144] (12 x 12 column-major).  Host pointers.
int qrw_solo12_feet_host(int batch, const double* q12, const double* dq12, double* posf, double* vf, double* wf,
                         double* af, double* Jf) {
  if (batch <= 0 || !q12 || !dq12 || !posf || !vf || !wf || !af || !Jf) return fail("qrw_solo12_feet_host: bad argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail("qrw_solo12_feet_host: no CUDA device (this library has no CPU fallback)");
  const size_t B = (size_t)batch;
  double* d = nullptr;
  CU(cudaMalloc(&d, B * (12 + 12 + 48 + 144) * sizeof(double)));
  double *dq_ = d + B * 12, *dp = d + B * 24, *dv = dp + B * 12, *dw = dv + B * 12, *da = dw + B * 12, *dJ = da + B * 12;
  cudaError_t e = cudaMemcpy(d, q12, B * 12 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(dq_, dq12, B * 12 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    solo12_feet_kernel<<<(unsigned)((4 * B + 127) / 128), 128>>>(batch, d, dq_, dp, dv, dw, da, dJ);
    ++g_launches;
    e = cudaGetLastError();
  }
  double* outs[4] = {posf, vf, wf, af};
  double* srcs[4] = {dp, dv, dw, da};
  for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaMemcpy(outs[i], srcs[i], B * 12 * sizeof(double), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(Jf, dJ, B * 144 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return fail(std::string("qrw_solo12_feet_host: ") + cudaGetErrorString(e));
  return 0;
}

}  // extern "C"
#include "planner_abi.cuh"
#include "tick_abi.cuh"
#include <time.h>
#include "transport_abi.cuh"
#include "noalloc_abi.cuh"
