// tools/tmem_bench.cu -- tensor memory (TMEM) as a per-lane scratchpad on B200: what tcgen05.ld / tcgen05.st cost when no
// tcgen05.mma is involved (not part of the product; evidence for DESIGN.md section 4.1).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tmem_bench.bin tools/tmem_bench.cu && tools/tmem_bench.bin
// Questions: (1) do two CTAs per SM co-reside when each allocates 256 columns; (2) latency of a dependent
// tcgen05.ld (32x32b: every thread reads consecutive columns of its own lane); (3) issue interval / bandwidth of
// back-to-back loads from 1, 4, 8 warps and from two co-resident CTAs; (4) does TMEM traffic delay shared-memory loads
// of another warp and the other way round (separate pipes?).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t tmem_alloc(uint32_t* slot, int ncols) {  // whole CTA calls it; warp 0 allocates
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *slot;
}
__device__ __forceinline__ void tmem_free(uint32_t base, int ncols) {
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(ncols) : "memory");
}
__device__ __forceinline__ void ld8(uint32_t a, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(a));
}
__device__ __forceinline__ void ld16(uint32_t a, uint32_t (&r)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(a));
}
__device__ __forceinline__ void st8(uint32_t a, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(a), "r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// (0) correctness + co-residency: every CTA allocates `ncols`, writes a pattern into its lanes, reads it back, and
// reports its SM and the interval (globaltimer) it was alive for
__global__ void k_alloc(int ncols, int spin, int* smid, unsigned long long* t0, unsigned long long* t1, int* bad) {
  extern __shared__ double pad[];
  __shared__ uint32_t slot;
  unsigned long long a, b;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(a));
  const uint32_t base = tmem_alloc(&slot, ncols);
  const int w = threadIdx.x >> 5;
  const uint32_t lane_base = base + ((uint32_t)(32 * (w & 3)) << 16);
  int nb = 0;
  // two warps share a quadrant: warp w uses the column half (w >> 2)
  const int c0 = (w >> 2) * (ncols / 2);
  for (int c = 0; c + 8 <= ncols / 2; c += 8) {
    uint32_t v[8];
    for (int q = 0; q < 8; ++q) v[q] = 1000003u * (blockIdx.x + 1) + 4099u * threadIdx.x + (c + q);
    st8(lane_base + c0 + c, v);
  }
  wait_st();
  for (volatile int s = 0; s < spin; ++s) {
  }
  for (int c = 0; c + 8 <= ncols / 2; c += 8) {
    uint32_t v[8];
    ld8(lane_base + c0 + c, v);
    wait_ld();
    for (int q = 0; q < 8; ++q) nb += v[q] != 1000003u * (blockIdx.x + 1) + 4099u * threadIdx.x + (c + q);
  }
  if (nb) atomicAdd(bad, nb);
  tmem_free(base, ncols);
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(b));
  if (threadIdx.x == 0) {
    int s;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(s));
    smid[blockIdx.x] = s, t0[blockIdx.x] = a, t1[blockIdx.x] = b;
  }
  pad[threadIdx.x] = 0.0;
}

// (1) dependent load: the address of load i+1 depends on a value of load i (the value is 0)
__global__ void k_ld_lat(int n, long long* cyc, uint32_t* out) {
  __shared__ uint32_t slot;
  const uint32_t base = tmem_alloc(&slot, 64);
  const uint32_t lb = base + ((uint32_t)(32 * ((threadIdx.x >> 5) & 3)) << 16);
  uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int c = 0; c < 64; c += 8) st8(lb + c, z);
  wait_st();
  uint32_t a = lb, acc = 0;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    uint32_t v[8];
    ld8(a, v);
    wait_ld();
    a = lb + v[0] + ((i & 3) << 3);
    acc += v[7];
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
  out[threadIdx.x] = acc + a;
  tmem_free(base, 64);
}
// (1b) store -> wait::st -> load -> wait::ld round trip, dependent through the value
__global__ void k_st_ld(int n, long long* cyc, uint32_t* out) {
  __shared__ uint32_t slot;
  const uint32_t base = tmem_alloc(&slot, 64);
  const uint32_t lb = base + ((uint32_t)(32 * ((threadIdx.x >> 5) & 3)) << 16);
  uint32_t v[8] = {1, 2, 3, 4, 5, 6, 7, 8};
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    st8(lb, v);
    wait_st();
    ld8(lb, v);
    wait_ld();
    v[0] += 1;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
  out[threadIdx.x] = v[0] + v[7];
  tmem_free(base, 64);
}
// (2) throughput: U independent x16 loads (24-column rows would be x16 + x8) back to back, one wait per batch
template <int U>
__global__ void k_ld_tput(int n, int ncols, long long* cyc, uint32_t* out) {
  __shared__ uint32_t slot;
  const uint32_t base = tmem_alloc(&slot, ncols);
  const int w = threadIdx.x >> 5;
  const uint32_t lb = base + ((uint32_t)(32 * (w & 3)) << 16) + (w >> 2) * (ncols / 2);
  uint32_t acc = 0;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    uint32_t v[U][16];
#pragma unroll
    for (int u = 0; u < U; ++u) ld16(lb + ((16 * u) & (ncols / 2 - 1)), v[u]);
    wait_ld();
#pragma unroll
    for (int u = 0; u < U; ++u) acc += v[u][0] ^ v[u][15];
  }
  long long t1 = clock64();
  if ((threadIdx.x & 31) == 0) cyc[blockIdx.x * 8 + w] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  tmem_free(base, ncols);
}
// (3) interference: warp 0 chases a pointer through shared memory (LDS latency) while warps 1..7 either idle (mode 0),
// stream tcgen05.ld (mode 1) or stream conflict-free LDS.128 (mode 2)
__global__ void k_mix(int n, int mode, long long* cyc, uint32_t* out) {
  __shared__ uint32_t slot;
  __shared__ int chase[1024];
  __shared__ double4 bulk[1024];
  const uint32_t base = tmem_alloc(&slot, 256);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) chase[i] = (i + 32) & 1023;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) bulk[i] = make_double4(i, 1, 2, 3);
  __shared__ volatile int stop;
  if (threadIdx.x == 0) stop = 0;
  __syncthreads();
  const uint32_t lb = base + ((uint32_t)(32 * (w & 3)) << 16) + (w >> 2) * 128;
  if (w == 0) {
    int p = l;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) p = chase[p];
    long long t1 = clock64();
    if (l == 0) cyc[0] = t1 - t0;
    out[l] = p;
    __syncwarp();
    if (l == 0) stop = 1;
  } else {
    uint32_t acc = 0;
    double s = 0.0;
    long long cnt = 0;
    while (!stop) {
      if (mode == 1) {
        uint32_t v[4][16];
#pragma unroll
        for (int u = 0; u < 4; ++u) ld16(lb + 16 * u, v[u]);
        wait_ld();
#pragma unroll
        for (int u = 0; u < 4; ++u) acc += v[u][3];
      } else if (mode == 2) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const double4 d = bulk[(l + 32 * u + 256 * (w - 1)) & 1023];
          s += d.x + d.w;
        }
      }
      ++cnt;
    }
    out[threadIdx.x] = acc + (uint32_t)s;
    if (l == 0) cyc[w] = cnt;
  }
  tmem_free(base, 256);
}
// (3b) the other way round: warp 0 runs dependent tcgen05.ld while the other warps stream LDS.128 (mode 2) or idle (0)
__global__ void k_mix2(int n, int mode, long long* cyc, uint32_t* out) {
  __shared__ uint32_t slot;
  __shared__ double4 bulk[1024];
  const uint32_t base = tmem_alloc(&slot, 256);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) bulk[i] = make_double4(i, 1, 2, 3);
  __shared__ volatile int stop;
  if (threadIdx.x == 0) stop = 0;
  const uint32_t lb = base + ((uint32_t)(32 * (w & 3)) << 16) + (w >> 2) * 128;
  uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int c = 0; c < 64; c += 8) st8(lb + c, z);
  wait_st();
  __syncthreads();
  if (w == 0) {
    uint32_t a = lb, acc = 0;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
      uint32_t v[8];
      ld8(a, v);
      wait_ld();
      a = lb + v[0] + ((i & 3) << 3);
      acc += v[7];
    }
    long long t1 = clock64();
    if (l == 0) cyc[0] = t1 - t0;
    out[l] = acc;
    __syncwarp();
    if (l == 0) stop = 1;
  } else {
    double s = 0.0;
    while (!stop) {
      if (mode == 2) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const double4 d = bulk[(l + 32 * u + 256 * (w - 1)) & 1023];
          s += d.x + d.w;
        }
      }
    }
    out[threadIdx.x] = (uint32_t)s;
  }
  tmem_free(base, 256);
}

int main() {
  cudaDeviceProp pr;
  cudaGetDeviceProperties(&pr, 0);
  const int nsm = pr.multiProcessorCount;
  printf("%s, %d SMs\n", pr.name, nsm);
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 22);
  cudaMallocManaged(&cyc, 1 << 16);
  // (0) co-residency
  for (int ncols : {128, 256, 512}) {
    const int grid = 2 * nsm;
    int *smid, *bad;
    unsigned long long *t0, *t1;
    cudaMallocManaged(&smid, grid * 4);
    cudaMallocManaged(&bad, 4);
    cudaMallocManaged(&t0, grid * 8);
    cudaMallocManaged(&t1, grid * 8);
    *bad = 0;
    cudaFuncSetAttribute(k_alloc, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    k_alloc<<<grid, 256, 100 * 1024>>>(ncols, 2000000, smid, t0, t1, bad);
    cudaError_t e = cudaDeviceSynchronize();
    int overlap = 0;
    for (int i = 0; i < grid; ++i)
      for (int j = i + 1; j < grid; ++j)
        if (smid[i] == smid[j]) {
          const unsigned long long lo = t0[i] > t0[j] ? t0[i] : t0[j], hi = t1[i] < t1[j] ? t1[i] : t1[j];
          if (hi > lo && (double)(hi - lo) > 0.5 * (double)(t1[i] - t0[i])) ++overlap;
        }
    printf("alloc %3d columns, 2 CTAs/SM by smem and registers: %d CTA pairs overlapped in time on one SM (of %d SMs), pattern errors %d, %s\n",
           ncols, overlap, nsm, *bad, cudaGetErrorString(e));
  }
  const int n = 4096;
  k_ld_lat<<<1, 32>>>(n, cyc, out); cudaDeviceSynchronize();
  printf("tcgen05.ld.32x32b.x8 + wait::ld, dependent (1 warp)        %8.2f cycles\n", (double)cyc[0] / n);
  k_st_ld<<<1, 32>>>(n, cyc, out); cudaDeviceSynchronize();
  printf("tcgen05.st + wait::st + tcgen05.ld + wait::ld (1 warp)      %8.2f cycles\n", (double)cyc[0] / n);
  auto tput = [&](int warps, int ctas_per_sm, const char* what) {
    const int grid = ctas_per_sm == 0 ? 1 : ctas_per_sm * nsm;
    k_ld_tput<8><<<grid, 32 * warps>>>(n, 256, cyc, out);
    cudaDeviceSynchronize();
    long long mx = 0;
    for (int b = 0; b < grid; ++b)
      for (int w = 0; w < warps; ++w) mx = cyc[b * 8 + w] > mx ? cyc[b * 8 + w] : mx;
    const double per = (double)mx / n / 8;  // cycles per x16 load per warp
    const int cps = ctas_per_sm == 0 ? 1 : ctas_per_sm;
    printf("tcgen05.ld.x16 stream, %-28s %6.2f cycles per load per warp -> %7.1f B/cycle/SM\n", what, per, 2048.0 * warps * cps / per);
  };
  tput(1, 0, "1 warp");
  tput(4, 0, "4 warps (4 quadrants)");
  tput(8, 0, "8 warps (2 per quadrant)");
  tput(8, 1, "8 warps, every SM busy");
  tput(8, 2, "8 warps x 2 CTAs/SM");
  for (int mode = 0; mode < 3; ++mode) {
    k_mix<<<1, 256>>>(n, mode, cyc, out); cudaDeviceSynchronize();
    printf("LDS dependent latency, other 7 warps %-22s %8.2f cycles\n", mode == 0 ? "idle" : mode == 1 ? "stream tcgen05.ld" : "stream LDS.128", (double)cyc[0] / n);
  }
  for (int mode = 0; mode < 3; mode += 2) {
    k_mix2<<<1, 256>>>(n, mode, cyc, out); cudaDeviceSynchronize();
    printf("tcgen05.ld dependent latency, other 7 warps %-15s %8.2f cycles\n", mode == 0 ? "idle" : "stream LDS.128", (double)cyc[0] / n);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
