// tools/fp64_bench.cu -- the FP64 roofs of this B200 (not part of the product): DFMA peak, fp64 tensor-core peak
// (mma.sync.aligned.m8n8k4.f64, "DMMA"; tcgen05.mma has no f64 kind) and cuBLAS DGEMM 8192^3.  SURVEY.md section 6 asks
// for these numbers before any factorisation roofline is quoted; DESIGN.md section 4.1 rests its "no tensor cores for
// the 12x12 factorisations" decision on them.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_bench.bin tools/fp64_bench.cu -lcublas && tools/fp64_bench.bin
#include <cstdio>
#include <cuda_runtime.h>
#include <cublas_v2.h>

// 8 independent DFMA chains per thread
__global__ void k_dfma(double* out, int n, double a, double b) {
  double v[8];
  for (int c = 0; c < 8; ++c) v[c] = threadIdx.x + c;
  for (int i = 0; i < n; ++i)
#pragma unroll
    for (int c = 0; c < 8; ++c) v[c] = fma(v[c], a, b);
  double s = 0;
  for (int c = 0; c < 8; ++c) s += v[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 8 independent DMMA accumulators per warp: D(8x8) += A(8x4) B(4x8), 2 accumulator doubles per thread
__global__ void k_dmma(double* out, int n, double a, double b) {
  double c[8][2];
  for (int q = 0; q < 8; ++q) c[q][0] = c[q][1] = threadIdx.x + q;
  for (int i = 0; i < n; ++i)
#pragma unroll
    for (int q = 0; q < 8; ++q)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[q][0]), "+d"(c[q][1]) : "d"(a), "d"(b));
  double s = 0;
  for (int q = 0; q < 8; ++q) s += c[q][0] + c[q][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// dependent DMMA chain: latency of one m8n8k4
__global__ void k_dmma_lat(double* out, long long* cyc, int n, double a, double b) {
  double c0 = threadIdx.x, c1 = 1.0;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i)
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
  long long t1 = clock64();
  out[threadIdx.x] = c0 + c1;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
  cudaDeviceProp pr;
  cudaGetDeviceProperties(&pr, 0);
  const int nsm = pr.multiProcessorCount;
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 24);
  cudaMallocManaged(&cyc, 64);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float ms;
  const int n = 1 << 15;
  {
    const int grid = nsm * 4, thr = 512;
    k_dfma<<<grid, thr>>>(out, 1024, 0.999, 1e-9);
    double best = 0;
    for (int r = 0; r < 5; ++r) {
      cudaEventRecord(e0);
      k_dfma<<<grid, thr>>>(out, n, 0.999, 1e-9);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1);
      const double tf = 2.0 * 8 * n * (double)grid * thr / (ms * 1e-3) / 1e12;
      best = tf > best ? tf : best;
    }
    printf("DFMA peak (8 chains/thread, %d x %d threads)        %8.2f TFLOP/s\n", grid, thr, best);
  }
  {
    const int grid = nsm * 4, thr = 512;
    k_dmma<<<grid, thr>>>(out, 1024, 0.999, 1e-9);
    double best = 0;
    for (int r = 0; r < 5; ++r) {
      cudaEventRecord(e0);
      k_dmma<<<grid, thr>>>(out, n, 0.999, 1e-9);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1);
      // one m8n8k4 = 8*8*4 FMA = 512 FLOP per warp
      const double tf = 512.0 * 8 * n * (double)grid * (thr / 32) / (ms * 1e-3) / 1e12;
      best = tf > best ? tf : best;
    }
    printf("DMMA peak (mma.sync m8n8k4 f64, 8 accumulators/warp)   %8.2f TFLOP/s\n", best);
    k_dmma_lat<<<1, 32>>>(out, cyc, 4096, 0.999, 1e-9);
    cudaDeviceSynchronize();
    printf("DMMA m8n8k4 dependent latency (1 warp)                 %8.2f cycles\n", (double)cyc[0] / 4096);
  }
  {
    const int N = 8192;
    double *A, *B, *C;
    cudaMalloc(&A, sizeof(double) * N * N);
    cudaMalloc(&B, sizeof(double) * N * N);
    cudaMalloc(&C, sizeof(double) * N * N);
    cudaMemset(A, 0, sizeof(double) * N * N);
    cudaMemset(B, 0, sizeof(double) * N * N);
    cublasHandle_t h;
    cublasCreate(&h);
    const double al = 1.0, be = 0.0;
    cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_N, N, N, N, &al, A, N, B, N, &be, C, N);
    double best = 0;
    for (int r = 0; r < 3; ++r) {
      cudaEventRecord(e0);
      cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_N, N, N, N, &al, A, N, B, N, &be, C, N);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1);
      const double tf = 2.0 * N * (double)N * N / (ms * 1e-3) / 1e12;
      best = tf > best ? tf : best;
    }
    printf("cuBLAS DGEMM 8192^3                                    %8.2f TFLOP/s\n", best);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
