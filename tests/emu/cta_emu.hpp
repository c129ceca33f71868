// tests/emu/cta_emu.hpp -- TEST INFRASTRUCTURE.  Runs the threads of one CUDA CTA as cooperative
// fibers on the calling OS thread so that the product's per-thread kernel code (csrc/mpc_cta_core.cuh,
// compiled with -DQRW_EMULATE) can be executed and compared with the oracle on a machine without a GPU.
// Fibers switch only at the collectives of qrw::cta (bar, wsync, shfl, warp_max, gather12); between two
// collectives a thread runs alone, which is a legal CUDA schedule.  x86-64 only (hand-written switch).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <vector>

#include "../../quadruped-reactive-walking_b200/csrc/qrw_spmd.cuh"  // declarations of the qrw::cta primitives defined here (TRow)

#if !defined(__x86_64__)
#error "cta_emu.hpp: x86-64 only"
#endif

namespace qrw {
namespace cta {

extern "C" void qrw_emu_switch(void** save_sp, void* load_sp);
asm(R"(
.text
.globl qrw_emu_switch
.type qrw_emu_switch,@function
qrw_emu_switch:
  pushq %rbp
  pushq %rbx
  pushq %r12
  pushq %r13
  pushq %r14
  pushq %r15
  movq %rsp, (%rdi)
  movq %rsi, %rsp
  popq %r15
  popq %r14
  popq %r13
  popq %r12
  popq %rbx
  popq %rbp
  ret
.size qrw_emu_switch,.-qrw_emu_switch
)");

struct Emu {
  enum { RUN = 0, AT_BAR = 1, AT_WSYNC = 2, DONE = 3, AT_NAMED = 4 };
  int nb_count[16] = {0};   // named barriers (bar.arrive / bar.sync id, count): threads seen so far
  std::vector<int> nb_id;   // id a thread in AT_NAMED waits on
  int n = 0, cur = -1;
  std::vector<void*> sp;
  std::vector<char*> stacks;
  std::vector<int> state;
  std::vector<double> xch;
  std::vector<double> tmem = std::vector<double>(128 * 128, 0.0);  // tensor memory of the CTA: 128 lanes x 256 columns (128 doubles)
  void* sched_sp = nullptr;
  std::function<void()> body;
  static constexpr size_t STACK = 256 * 1024;
};
inline thread_local Emu* g_emu = nullptr;
// order in which the scheduler sweeps the runnable fibers: 0 ascending thread id, 1 descending, 2 odd warps first (lanes descending),
// 3 warps descending, 4 lanes descending.
// Every order is a legal CUDA schedule; code without races between collectives gives bit-identical results in all of them.
inline int g_sweep_order = 0;

inline void fiber_entry() {
  Emu* e = g_emu;
  e->body();
  e->state[e->cur] = Emu::DONE;
  qrw_emu_switch(&e->sp[e->cur], e->sched_sp);
  abort();
}
inline void yield_as(int st) {
  Emu* e = g_emu;
  e->state[e->cur] = st;
  qrw_emu_switch(&e->sp[e->cur], e->sched_sp);
}

// Run `body` once per thread of an `nthreads`-wide CTA.
inline void run_cta(int nthreads, std::function<void()> body) {
  Emu e;
  e.n = nthreads;
  e.body = std::move(body);
  e.sp.resize(nthreads);
  e.stacks.resize(nthreads);
  e.state.assign(nthreads, Emu::RUN);
  e.xch.assign(nthreads, 0.0);
  e.nb_id.assign(nthreads, -1);
  for (int t = 0; t < nthreads; ++t) {
    char* st = (char*)aligned_alloc(64, Emu::STACK);
    e.stacks[t] = st;
    uintptr_t top = ((uintptr_t)(st + Emu::STACK)) & ~(uintptr_t)15;
    void** p = (void**)top;
    *--p = nullptr;              // fake return address of fiber_entry
    *--p = (void*)&fiber_entry;  // `ret` target
    for (int r = 0; r < 6; ++r) *--p = nullptr;
    e.sp[t] = (void*)p;
  }
  Emu* prev = g_emu;
  g_emu = &e;
  for (;;) {
    bool progress = false;
    int alive = 0;
    for (int tt = 0; tt < nthreads; ++tt) {
      int t = tt;
      if (g_sweep_order == 1) t = nthreads - 1 - tt;
      else if (g_sweep_order == 3) t = (nthreads / 32 - 1 - tt / 32) * 32 + (tt & 31);  // warps descending, lanes ascending
      else if (g_sweep_order == 4) t = (tt & ~31) + 31 - (tt & 31);                       // warps ascending, lanes descending
      else if (g_sweep_order == 2 && nthreads > 32) {
        const int nw = nthreads / 32, w = tt / 32;
        const int wsel = w < nw / 2 ? 2 * w + 1 : 2 * (w - nw / 2);  // 1, 3, 5, .., 0, 2, 4, ..
        t = (wsel < nw ? wsel : w) * 32 + 31 - (tt & 31);
      }
      if (e.state[t] == Emu::RUN) {
        e.cur = t;
        qrw_emu_switch(&e.sched_sp, e.sp[t]);
        progress = true;
      }
      if (e.state[t] != Emu::DONE) ++alive;
    }
    if (!alive) break;
    // release barriers whose participants have all arrived
    bool released = false;
    bool all_bar = true;
    for (int t = 0; t < nthreads; ++t)
      if (e.state[t] != Emu::AT_BAR && e.state[t] != Emu::DONE) all_bar = false;
    if (all_bar) {
      for (int t = 0; t < nthreads; ++t)
        if (e.state[t] == Emu::AT_BAR) e.state[t] = Emu::RUN;
      released = true;
    }
    for (int w0 = 0; w0 < nthreads; w0 += 32) {
      bool all = true, any = false;
      for (int t = w0; t < w0 + 32 && t < nthreads; ++t) {
        if (e.state[t] == Emu::AT_WSYNC) any = true;
        else if (e.state[t] != Emu::DONE) all = false;
      }
      if (all && any) {
        for (int t = w0; t < w0 + 32 && t < nthreads; ++t)
          if (e.state[t] == Emu::AT_WSYNC) e.state[t] = Emu::RUN;
        released = true;
      }
    }
    if (!progress && !released) {
      fprintf(stderr, "cta_emu: deadlock (divergent barrier)\n");
      abort();
    }
  }
  g_emu = prev;
  for (char* s : e.stacks) free(s);
}

inline int tid() { return g_emu->cur; }
inline void bar() { yield_as(Emu::AT_BAR); }
inline void wsync() { yield_as(Emu::AT_WSYNC); }
// named barrier `id` completes when `nthreads` threads have arrived (bar.arrive: without waiting, bar.sync: waiting)
inline bool nb_count_in(int id, int nthreads) {
  Emu* e = g_emu;
  if (++e->nb_count[id] < nthreads) return false;
  e->nb_count[id] = 0;
  for (int t = 0; t < e->n; ++t)
    if (e->state[t] == Emu::AT_NAMED && e->nb_id[t] == id) e->state[t] = Emu::RUN;
  return true;
}
inline void nb_arrive(int id, int nthreads) { nb_count_in(id, nthreads); }
inline void nb_sync(int id, int nthreads) {
  if (nb_count_in(id, nthreads)) return;
  g_emu->nb_id[g_emu->cur] = id;
  yield_as(Emu::AT_NAMED);
}
inline double shfl(double v, int src) {
  Emu* e = g_emu;
  const int t = e->cur;
  e->xch[t] = v;
  wsync();
  const double r = e->xch[(t & ~31) + (src & 31)];
  wsync();
  return r;
}
inline double warp_max(double v) {
  Emu* e = g_emu;
  const int t = e->cur, w0 = t & ~31;
  e->xch[t] = v;
  wsync();
  double m = e->xch[w0];
  for (int i = 1; i < 32 && w0 + i < e->n; ++i) m = e->xch[w0 + i] > m ? e->xch[w0 + i] : m;
  wsync();
  return m;
}
inline void gather12(double v, double* out) {
  Emu* e = g_emu;
  const int t = e->cur, w0 = t & ~31;
  e->xch[t] = v;
  wsync();
  for (int j = 0; j < 12; ++j) out[j] = e->xch[w0 + j];
  wsync();
}
inline bool any(bool p) {
  Emu* e = g_emu;
  const int t = e->cur, w0 = t & ~31;
  e->xch[t] = p ? 1.0 : 0.0;
  wsync();
  bool r = false;
  for (int i = 0; i < 32 && w0 + i < e->n; ++i) r = r || e->xch[w0 + i] != 0.0;
  wsync();
  return r;
}
inline double shfl16(double v, int src) {
  Emu* e = g_emu;
  const int t = e->cur;
  e->xch[t] = v;
  wsync();
  const double r = e->xch[(t & ~15) + (src & 15)];
  wsync();
  return r;
}
inline double shfl_xor(double v, int m) {
  Emu* e = g_emu;
  const int t = e->cur;
  e->xch[t] = v;
  wsync();
  const double r = e->xch[(t & ~31) + ((t ^ m) & 31)];
  wsync();
  return r;
}
// tensor memory: a thread reaches the lane (warp % 4) * 32 + lane of its quadrant and nothing else
inline unsigned tmem_addr(unsigned base, int col) { return base + ((unsigned)(g_emu->cur & 96) << 16) + (unsigned)col; }
inline double* tmem_row(unsigned addr, int ncol = 24) {
  Emu* e = g_emu;
  const int lane0 = (int)(addr >> 16), col = (int)(addr & 0xffff);
  if (lane0 != (e->cur & 96) || (col & 1) || col + ncol > 256) {
    fprintf(stderr, "cta_emu: bad TMEM access by thread %d (lane base %d, column %d)\n", e->cur, lane0, col);
    abort();
  }
  return e->tmem.data() + (size_t)(lane0 + (e->cur & 31)) * 128 + col / 2;
}
inline void tmem_ld12(unsigned addr, TRow& t) {
  const double* p = tmem_row(addr);
  for (int i = 0; i < 12; ++i) t.d[i] = p[i];
}
inline void tmem_st12(unsigned addr, const double* v) {
  double* p = tmem_row(addr);
  for (int i = 0; i < 12; ++i) p[i] = v[i];
}
inline void tmem_ld4(unsigned addr, TRow4& t) {
  const double* p = tmem_row(addr, 8);
  for (int i = 0; i < 4; ++i) t.d[i] = p[i];
}
inline void tmem_st4(unsigned addr, const double* v) {
  double* p = tmem_row(addr, 8);
  for (int i = 0; i < 4; ++i) p[i] = v[i];
}
inline int fetch_add(int* p, int v) {
  const int o = *p;
  *p = o + v;
  return o;
}

}  // namespace cta
}  // namespace qrw
