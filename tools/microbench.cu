// tools/microbench.cu -- latencies that bound the MPC kernel's critical path on B200 (not part of the product).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/microbench tools/microbench.cu && /tmp/microbench
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_dfma_chain(double* out, long long* cyc, int n, double a, double b) {
  double v = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) v = fma(v, a, b);
  long long t1 = clock64();
  out[threadIdx.x] = v;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_dadd_chain(double* out, long long* cyc, int n, double b) {
  double v = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) v = v + b;
  long long t1 = clock64();
  out[threadIdx.x] = v;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
// NCH independent chains per thread: issue rate of DFMA for one warp
template <int NCH>
__global__ void k_dfma_ilp(double* out, long long* cyc, int n, double a, double b) {
  double v[NCH];
  for (int c = 0; c < NCH; ++c) v[c] = threadIdx.x + c;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i)
#pragma unroll
    for (int c = 0; c < NCH; ++c) v[c] = fma(v[c], a, b);
  long long t1 = clock64();
  double s = 0;
  for (int c = 0; c < NCH; ++c) s += v[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
// pointer chase in shared memory: LDS latency
__global__ void k_lds_chain(int* out, long long* cyc, int n) {
  __shared__ int s[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) s[i] = (i + 32) & 1023;
  __syncthreads();
  int p = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) p = s[p];
  long long t1 = clock64();
  out[threadIdx.x] = p;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
// STS -> __syncwarp -> LDS (another lane's value) round trip, dependent
__global__ void k_sts_lds(double* out, long long* cyc, int n) {
  __shared__ double s[64];
  double v = threadIdx.x;
  const int src = (threadIdx.x + 1) & 31;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    s[threadIdx.x] = v;
    __syncwarp();
    v = s[src] + 1.0;
    __syncwarp();
  }
  long long t1 = clock64();
  out[threadIdx.x] = v;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_shfl(double* out, long long* cyc, int n) {
  double v = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) v = __shfl_sync(0xffffffffu, v, (threadIdx.x + 1) & 31) + 1.0;
  long long t1 = clock64();
  out[threadIdx.x] = v;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_bar(double* out, long long* cyc, int n) {
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) __syncthreads();
  long long t1 = clock64();
  out[threadIdx.x] = 0;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMallocManaged(&cyc, 4096 * sizeof(long long));
  const int n = 4096;
  auto rep = [&](const char* name, double per) { printf("%-44s %8.2f cycles\n", name, per); };
  k_dfma_chain<<<1, 32>>>(out, cyc, n, 0.999, 1e-9); cudaDeviceSynchronize(); rep("DFMA dependent latency (1 warp)", (double)cyc[0] / n);
  k_dadd_chain<<<1, 32>>>(out, cyc, n, 1e-9); cudaDeviceSynchronize(); rep("DADD dependent latency (1 warp)", (double)cyc[0] / n);
  k_dfma_ilp<4><<<1, 32>>>(out, cyc, n, 0.999, 1e-9); cudaDeviceSynchronize(); rep("DFMA issue interval, 1 warp, 4 chains", (double)cyc[0] / n / 4);
  k_dfma_ilp<8><<<1, 32>>>(out, cyc, n, 0.999, 1e-9); cudaDeviceSynchronize(); rep("DFMA issue interval, 1 warp, 8 chains", (double)cyc[0] / n / 8);
  k_dfma_ilp<8><<<1, 128>>>(out, cyc, n, 0.999, 1e-9); cudaDeviceSynchronize(); rep("DFMA issue interval per SMSP, 4 warps x 8 chains", (double)cyc[0] / n / 8);
  k_dfma_ilp<8><<<1, 512>>>(out, cyc, n, 0.999, 1e-9); cudaDeviceSynchronize(); rep("DFMA issue interval per SMSP, 16 warps (4/SMSP)", (double)cyc[0] / n / 8 / 4);
  k_dfma_ilp<1><<<1, 512>>>(out, cyc, n, 0.999, 1e-9); cudaDeviceSynchronize(); rep("DFMA chain step with 4 warps/SMSP, 1 chain each", (double)cyc[0] / n);
  k_lds_chain<<<1, 32>>>((int*)out, cyc, n); cudaDeviceSynchronize(); rep("LDS dependent latency", (double)cyc[0] / n);
  k_sts_lds<<<1, 32>>>(out, cyc, n); cudaDeviceSynchronize(); rep("STS + syncwarp + LDS + DADD + syncwarp", (double)cyc[0] / n);
  k_shfl<<<1, 32>>>(out, cyc, n); cudaDeviceSynchronize(); rep("SHFL(double) + DADD dependent", (double)cyc[0] / n);
  k_bar<<<1, 256>>>(out, cyc, n); cudaDeviceSynchronize(); rep("__syncthreads, 8 warps", (double)cyc[0] / n);
  k_bar<<<1, 32>>>(out, cyc, n); cudaDeviceSynchronize(); rep("__syncthreads, 1 warp", (double)cyc[0] / n);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
