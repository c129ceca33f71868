// tools/chain_bench.cu -- the MPC kernel's state chain (qrw::v2::state_recurrence) alone on one SM, cycles per call and per
// segment (not part of the product).  Build twice:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/chain_bench.bin tools/chain_bench.cu
//   nvcc ... -DQRW_V2_TIMING -o tools/chain_bench_t.bin tools/chain_bench.cu      (adds the per-segment stamps)
#include <cstdio>
#include <cuda_runtime.h>
#include "../quadruped-reactive-walking_b200/csrc/mpc_v2_core.cuh"

using namespace qrw::v2;

__global__ void __launch_bounds__(256, 2) k_chain(double* out, long long* cyc, int n, int nwarp_active) {
  extern __shared__ double slab[];
  const int tid = threadIdx.x;
  for (int i = tid; i < V_W; i += blockDim.x) {
    double v = 0.001 * ((i * 2654435761u) % 1000) - 0.5;
    slab[i] = i < V_REC ? 0.02 * v : v;
  }
  __syncthreads();
  long long t0 = clock64();
  long long seg[4] = {0, 0, 0, 0};
  for (int it = 0; it < n; ++it) {
    if ((tid >> 5) == 0) state_recurrence(slab + V_SIG, slab + V_T, slab + V_REC, tid & 31);
    else if ((tid >> 5) == 1) state_helper(slab + V_SIG, slab + V_REC, tid & 31);
#ifdef QRW_V2_TIMING
    __syncwarp();
    if (tid == 0)
      for (int q = 0; q < 4; ++q) seg[q] += ((volatile long long*)chain_stamp)[q + 1] - ((volatile long long*)chain_stamp)[q];
#endif
    __syncthreads();
  }
  long long t1 = clock64();
  if (tid == 0) {
    cyc[0] = t1 - t0;
    for (int q = 0; q < 4; ++q) cyc[1 + q] = seg[q];
  }
  if (tid < 192) out[tid] = slab[V_REC + SR * (tid / 12) + R_XT + tid % 12];
}

int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMallocManaged(&cyc, 64 * sizeof(long long));
  const int n = 2000;
  const int smem = V_W * 8;
  cudaFuncSetAttribute(k_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int threads : {64, 256}) {
    k_chain<<<1, threads, smem>>>(out, cyc, n, 1);
    cudaDeviceSynchronize();
    printf("chain alone, CTA of %3d threads: %8.1f cycles per call (incl. one __syncthreads)", threads, (double)cyc[0] / n);
#ifdef QRW_V2_TIMING
    printf(" | elimination %.0f middle %.0f v-products %.0f back-substitution %.0f", (double)cyc[1] / n, (double)cyc[2] / n, (double)cyc[3] / n,
           (double)cyc[4] / n);
#endif
    printf("\n");
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
