// qrw_spmd.cuh -- warp-SPMD portability layer.
//
// The solver cores (mpc_core.cuh, wbc_core.cuh) are written as a sequence of "phases": a phase
// body runs once per lane of ONE warp, phases are separated by warp barriers, and all
// cross-lane traffic goes through the warp's shared-memory slab.  Under nvcc a phase body is
// executed by the 32 hardware lanes; under a plain C++ compiler (tests/emu, -DQRW_EMULATE) the
// same source runs the 32 lanes one after the other, which lets the kernel logic be checked
// against the CPU oracle on a machine without a GPU.  The emulation build is test
// infrastructure only -- the shipped library contains the CUDA build and nothing else.
//
// Rules that keep both builds equivalent:
//   * inside one phase a lane never reads a shared location another lane writes in that phase;
//   * code between phases ("uniform code") only reads shared memory (or writes values every
//     lane agrees on) and computes lane-invariant scalars.
#pragma once

#if defined(__CUDACC__) && !defined(QRW_EMULATE)
#define QRW_HD __device__ __forceinline__
#define QRW_HDN __device__ __noinline__
#define QRW_LANES(l) for (int l = (int)(threadIdx.x & 31), _once = 1; _once; _once = 0)
#define QRW_SYNC() __syncwarp()
#define QRW_ON_GPU 1
#else
#define QRW_HD inline
#define QRW_HDN inline
// lanes of a phase one after the other, ascending or (qrw::g_emu_lanes_descending, set by the tests) descending: a phase that
// keeps the rules below gives the same bits either way
namespace qrw {
inline int g_emu_lanes_descending = 0;
}
#define QRW_LANES(l) for (int _i = 0, l = qrw::g_emu_lanes_descending ? 31 : 0; _i < 32; ++_i, l += qrw::g_emu_lanes_descending ? -1 : 1)
#define QRW_SYNC() ((void)0)
#define QRW_ON_GPU 0
#include <cmath>
#endif

namespace qrw {

QRW_HD double dmax(double a, double b) { return a > b ? a : b; }
QRW_HD double dmin(double a, double b) { return a < b ? a : b; }

// OSQP limit_scaling (scaling.c): MIN_SCALING 1e-4 -> 1, MAX_SCALING 1e4
QRW_HD double limit_scaling(double v) {
  v = v < 1e-4 ? 1.0 : v;
  return v > 1e4 ? 1e4 : v;
}

// max over the 32 per-lane partials red[0..31] (uniform code; callers sync before and after)
QRW_HD double red_max32(const double* red) {
  double m = red[0];
#pragma unroll
  for (int i = 1; i < 32; ++i) m = dmax(m, red[i]);
  return m;
}
QRW_HD double red_sum32(const double* red) {
  double m = red[0];
#pragma unroll
  for (int i = 1; i < 32; ++i) m += red[i];
  return m;
}

}  // namespace qrw

// ------------------------------------------------------------------------------------------
// CTA-SPMD layer (mpc_cta_core.cuh): code written per THREAD of a multi-warp CTA, with the few
// collective primitives below.  Under nvcc they are the hardware ones; under -DQRW_EMULATE they
// are provided by tests/emu/cta_emu.hpp, which runs every thread of the CTA as a fiber and
// switches at the collectives (test infrastructure only).
// ------------------------------------------------------------------------------------------
#if QRW_ON_GPU
namespace qrw {
namespace cta {
__device__ __forceinline__ int tid() { return (int)threadIdx.x; }
__device__ __forceinline__ void bar() { __syncthreads(); }
__device__ __forceinline__ void wsync() { __syncwarp(); }
// 64-bit shuffles as two 32-bit ones on an explicitly unpacked pair: the built-in double overload leaves ptxas with the halves
// in swapped registers and three XORs per shuffle to put them back (seen in the SASS of the whole-body tick kernel)
__device__ __forceinline__ double shfl_idx64(double v, int src, int clamp) {
  double r;
  asm volatile(
      "{ .reg .b32 lo, hi;\n\t"
      "mov.b64 {lo, hi}, %1;\n\t"
      "shfl.sync.idx.b32 lo, lo, %2, %3, 0xffffffff;\n\t"
      "shfl.sync.idx.b32 hi, hi, %2, %3, 0xffffffff;\n\t"
      "mov.b64 %0, {lo, hi}; }"
      : "=d"(r)
      : "d"(v), "r"(src), "r"(clamp));
  return r;
}
__device__ __forceinline__ double shfl_bfly64(double v, int m) {
  double r;
  asm volatile(
      "{ .reg .b32 lo, hi;\n\t"
      "mov.b64 {lo, hi}, %1;\n\t"
      "shfl.sync.bfly.b32 lo, lo, %2, 0x1f, 0xffffffff;\n\t"
      "shfl.sync.bfly.b32 hi, hi, %2, 0x1f, 0xffffffff;\n\t"
      "mov.b64 %0, {lo, hi}; }"
      : "=d"(r)
      : "d"(v), "r"(m));
  return r;
}
__device__ __forceinline__ double shfl(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
// max over the warp, result on every lane
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, m));
  return v;
}
// out[j] = value held by lane j, j < 12, on every lane
__device__ __forceinline__ void gather12(double v, double* out) {
#pragma unroll
  for (int j = 0; j < 12; ++j) out[j] = __shfl_sync(0xffffffffu, v, j);
}
__device__ __forceinline__ int fetch_add(int* p, int v) { return atomicAdd(p, v); }
// true on every lane of the warp if the predicate holds on any of them
__device__ __forceinline__ bool any(bool p) { return __any_sync(0xffffffffu, p) != 0; }
// value held by lane `src` (0..15) of the caller's 16-lane group
__device__ __forceinline__ double shfl16(double v, int src) { return __shfl_sync(0xffffffffu, v, src, 16); }
__device__ __forceinline__ double shfl_xor(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
// warp_max / shfl on the explicitly unpacked pair (the MPC residual check: 72 shuffles per check)
__device__ __forceinline__ double warp_max_x(double v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v = fmax(v, shfl_bfly64(v, m));
  return v;
}
__device__ __forceinline__ double bcast(double v, int src) { return shfl_idx64(v, src, 0x1f); }
// the same two on the explicitly unpacked pair (whole-body tick)
__device__ __forceinline__ double hshfl(double v, int src) { return shfl_idx64(v, src, 0x101f); }
__device__ __forceinline__ double hxor(double v, int m) { return shfl_bfly64(v, m); }
// v[0..11] = p[0..11] from shared memory (16-byte aligned) as six LDS.128 issued back to back: volatile, so
// that ptxas keeps them together ahead of their consumers instead of serialising load -> FMA -> load.
__device__ __forceinline__ void lds12(const double* p, double* v) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(p);
#pragma unroll
  for (int i = 0; i < 12; i += 2)
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v[i]), "=d"(v[i + 1]) : "r"(a + 8 * i));
}
// v[i] = p[STRIDE * i], i = 0..11 (a column of a padded 12x12 block) as twelve LDS.64 that stay where they are written
template <int STRIDE>
__device__ __forceinline__ void lds12c(const double* p, double* v) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(p);
#pragma unroll
  for (int i = 0; i < 12; ++i) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v[i]) : "r"(a + 8 * STRIDE * i));
}
// v[0..5] = p[0..5] (16-byte aligned) as three LDS.128 issued back to back
__device__ __forceinline__ void lds6(const double* p, double* v) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(p);
#pragma unroll
  for (int i = 0; i < 6; i += 2)
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v[i]), "=d"(v[i + 1]) : "r"(a + 8 * i));
}
// named barriers of the CTA (ids 1..15; 0 is __syncthreads): `nthreads` = arriving + waiting threads, a multiple of 32
__device__ __forceinline__ void nb_arrive(int id, int nthreads) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }
__device__ __forceinline__ void nb_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }
// the value becomes unknown to the compiler (no instruction is emitted)
__device__ __forceinline__ void opaque(int& v) { asm volatile("" : "+r"(v)); }
// Scheduling fence for ptxas: a warp barrier is never crossed by its instruction scheduler, so the loads issued
// before it stay grouped ahead of the arithmetic that follows (costs one issue slot).
__device__ __forceinline__ void sched_fence() { __syncwarp(); }
// ------------------------------------------------------------------------------------------
// Tensor memory (TMEM, 256 KB per SM next to the tensor cores) used as a PER-LANE scratchpad for matrix rows -- no
// tcgen05.mma involved.  tcgen05.ld/st.32x32b: thread t of warp w reads/writes consecutive 32-bit columns of TMEM lane
// 32 (w % 4) + t, so a row a lane reads again in every ADMM iteration can live there instead of in shared memory.
// Measured on B200 (tools/tmem_bench.cu, profiles/r02_tmem_bench.txt): dependent load 34 cycles (LDS 29), 840 B/cycle/SM
// (shared memory 128), and -- the point -- TMEM loads neither queue behind shared-memory traffic nor delay it
// (LDS latency under an LDS.128 stream of the other warps: 163 cycles; under a tcgen05.ld stream: 30).
// Addresses: bits 31..16 lane, bits 15..0 column.  Everything here is warp-collective (.sync.aligned): call it from
// warp-uniform code only.  A row = 12 doubles = 24 columns = one x16 + one x8 access.
// ------------------------------------------------------------------------------------------
constexpr int TMEM_COLS = 256;  // columns a CTA allocates: two CTAs per SM share the 512
// whole CTA calls it (it ends with a CTA barrier); warp 0 allocates.  Returns the base address.
__device__ __forceinline__ unsigned tmem_alloc(unsigned* slot) {
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *slot;
}
__device__ __forceinline__ void tmem_free(unsigned base) {
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(TMEM_COLS) : "memory");
}
// address of column `col` in the calling warp's lane quadrant
__device__ __forceinline__ unsigned tmem_addr(unsigned base, int col) { return base + ((unsigned)(threadIdx.x & 96) << 16) + (unsigned)col; }
struct TRow {
  unsigned r[24];
};
// request the 12 doubles at `addr` (asynchronous; tmem_get12 waits)
__device__ __forceinline__ void tmem_ld12(unsigned addr, TRow& t) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(t.r[0]), "=r"(t.r[1]), "=r"(t.r[2]), "=r"(t.r[3]), "=r"(t.r[4]), "=r"(t.r[5]), "=r"(t.r[6]), "=r"(t.r[7]),
                 "=r"(t.r[8]), "=r"(t.r[9]), "=r"(t.r[10]), "=r"(t.r[11]), "=r"(t.r[12]), "=r"(t.r[13]), "=r"(t.r[14]), "=r"(t.r[15])
               : "r"(addr));
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(t.r[16]), "=r"(t.r[17]), "=r"(t.r[18]), "=r"(t.r[19]), "=r"(t.r[20]), "=r"(t.r[21]), "=r"(t.r[22]), "=r"(t.r[23])
               : "r"(addr + 16));
}
// wait for every outstanding tcgen05.ld of the thread, then hand the row out as doubles.  The registers are in/out
// operands of the wait so that the compiler cannot move a use in front of it.
__device__ __forceinline__ void tmem_get12(TRow& t, double* v) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(t.r[0]), "+r"(t.r[1]), "+r"(t.r[2]), "+r"(t.r[3]), "+r"(t.r[4]), "+r"(t.r[5]), "+r"(t.r[6]), "+r"(t.r[7]),
                 "+r"(t.r[8]), "+r"(t.r[9]), "+r"(t.r[10]), "+r"(t.r[11]), "+r"(t.r[12]), "+r"(t.r[13]), "+r"(t.r[14]), "+r"(t.r[15]),
                 "+r"(t.r[16]), "+r"(t.r[17]), "+r"(t.r[18]), "+r"(t.r[19]), "+r"(t.r[20]), "+r"(t.r[21]), "+r"(t.r[22]), "+r"(t.r[23])
               :
               : "memory");
#pragma unroll
  for (int i = 0; i < 12; ++i) v[i] = __hiloint2double((int)t.r[2 * i + 1], (int)t.r[2 * i]);
}
__device__ __forceinline__ void tmem_st12(unsigned addr, const double* v) {
  unsigned r[24];
#pragma unroll
  for (int i = 0; i < 12; ++i) r[2 * i] = (unsigned)__double2loint(v[i]), r[2 * i + 1] = (unsigned)__double2hiint(v[i]);
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(addr),
               "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
               "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
               : "memory");
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(addr + 16), "r"(r[16]), "r"(r[17]),
               "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23])
               : "memory");
}
// four doubles (8 columns): the tail of a 28-entry row segment
struct TRow4 {
  unsigned r[8];
};
__device__ __forceinline__ void tmem_ld4(unsigned addr, TRow4& t) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(t.r[0]), "=r"(t.r[1]), "=r"(t.r[2]), "=r"(t.r[3]), "=r"(t.r[4]), "=r"(t.r[5]), "=r"(t.r[6]), "=r"(t.r[7])
               : "r"(addr));
}
__device__ __forceinline__ void tmem_get4(TRow4& t, double* v) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(t.r[0]), "+r"(t.r[1]), "+r"(t.r[2]), "+r"(t.r[3]), "+r"(t.r[4]), "+r"(t.r[5]), "+r"(t.r[6]), "+r"(t.r[7])
               :
               : "memory");
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __hiloint2double((int)t.r[2 * i + 1], (int)t.r[2 * i]);
}
__device__ __forceinline__ void tmem_st4(unsigned addr, const double* v) {
  unsigned r[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) r[2 * i] = (unsigned)__double2loint(v[i]), r[2 * i + 1] = (unsigned)__double2hiint(v[i]);
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
// v[0..3] = p[0..3] (16-byte aligned) as two LDS.128
__device__ __forceinline__ void lds4(const double* p, double* v) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(p);
#pragma unroll
  for (int i = 0; i < 4; i += 2)
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v[i]), "=d"(v[i + 1]) : "r"(a + 8 * i));
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
}  // namespace cta
}  // namespace qrw
#else
namespace qrw {
namespace cta {  // implemented in tests/emu/cta_emu.hpp
int tid();
void bar();
void wsync();
double shfl(double v, int src);
double warp_max(double v);
void gather12(double v, double* out);
int fetch_add(int* p, int v);
bool any(bool p);
double shfl16(double v, int src);
double shfl_xor(double v, int m);
inline double warp_max_x(double v) { return warp_max(v); }
inline double bcast(double v, int src) { return shfl(v, src); }
inline double hshfl(double v, int src) { return shfl16(v, src); }
inline double hxor(double v, int m) { return shfl_xor(v, m); }
inline void sched_fence() {}
inline void opaque(int&) {}
void nb_arrive(int id, int nthreads);
void nb_sync(int id, int nthreads);
inline void lds12(const double* p, double* v) {
  for (int i = 0; i < 12; ++i) v[i] = p[i];
}
inline void lds6(const double* p, double* v) {
  for (int i = 0; i < 6; ++i) v[i] = p[i];
}
template <int STRIDE>
inline void lds12c(const double* p, double* v) {
  for (int i = 0; i < 12; ++i) v[i] = p[STRIDE * i];
}
// tensor-memory scratchpad: tests/emu/cta_emu.hpp keeps one array per emulated CTA, private to (lane quadrant, lane)
constexpr int TMEM_COLS = 256;
struct TRow {
  double d[12];
};
unsigned tmem_addr(unsigned base, int col);
void tmem_ld12(unsigned addr, TRow& t);
inline void tmem_get12(TRow& t, double* v) {
  for (int i = 0; i < 12; ++i) v[i] = t.d[i];
}
void tmem_st12(unsigned addr, const double* v);
struct TRow4 {
  double d[4];
};
void tmem_ld4(unsigned addr, TRow4& t);
inline void tmem_get4(TRow4& t, double* v) {
  for (int i = 0; i < 4; ++i) v[i] = t.d[i];
}
void tmem_st4(unsigned addr, const double* v);
inline void lds4(const double* p, double* v) {
  for (int i = 0; i < 4; ++i) v[i] = p[i];
}
inline void tmem_wait_st() {}
}  // namespace cta
}  // namespace qrw
#endif
