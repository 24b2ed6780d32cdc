// sim_runtime.h — the "GPU" of the protocol simulator.
//  * one OS thread per CTA; the CUDA threads of a CTA are fibers (ucontext) scheduled round-robin on that OS thread, and
//    __syncthreads() is "yield until every fiber of the CTA has arrived". ThreadSanitizer therefore sees one thread per
//    CTA: it checks exactly what crosses CTAs and ranks (the flag protocols), reliably — with one OS thread per CUDA thread
//    a run has thousands of threads, more than the sanitizer's 256 thread slots, and it starts to forget accesses.
//    (Build with --param=tsan-instrument-func-entry-exit=0: the sanitizer must not keep a shadow call stack across fiber
//    switches.)
//  * every rank owns a slice of one contiguous heap (rank p at heap + p * bytes, which is what the VMM mapping gives on the
//    real node) and issues its launches from its own OS thread, one after the other, like a stream.
#pragma once
#include <pthread.h>
#include <ucontext.h>

#include <algorithm>
#include <atomic>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <vector>

#include "kernels/common.cuh"  // the REAL csrc/kernels/common.cuh (+ device/symm.cuh), staged next to the host ptx.cuh

thread_local sim_dim3 threadIdx, blockIdx, blockDim, gridDim;
thread_local SimCta* sim_cta = nullptr;
thread_local unsigned sim_rng = 1;
thread_local int sim_rank = 0;
SimMc sim_mc;
namespace tdist {
void set_error(const std::string& msg) { fprintf(stderr, "tdist error: %s\n", msg.c_str()); }
}  // namespace tdist

namespace sim {

constexpr size_t kFiberStack = 64 << 10;

struct Fiber {
  ucontext_t ctx;
  std::unique_ptr<char[]> stack;
  bool done = false, at_barrier = false, at_warp = false;
  unsigned warp_mask = 0xffffffffu;  // lanes named by the pending __syncwarp / shuffle
};
struct CtaRun {
  const std::function<void()>* kernel;
  int rank, cta, ctas, threads;
  int gx = 0, gy = 1, gz = 1;  // 3-D grid extents (gx == 0: 1-D launch of `ctas` CTAs)
  SimCta state;
  ucontext_t sched;
  std::vector<Fiber> fibers;
  std::vector<char> warp_scratch;  // per warp: 32 x 16 bytes for shuffles, 32 x 32 bytes for the matrix instructions
  int cur = 0;
};
inline thread_local CtaRun* tl_cta = nullptr;

// TDIST_SIM_WATCHDOG_S=n: after n seconds print, for every CTA that is still running, how many of its threads have finished,
// wait at the CTA barrier, wait at a warp barrier or are runnable (polling) — then abort. For debugging a scenario that hangs.
inline std::mutex g_live_mu;
inline std::vector<CtaRun*> g_live;
inline void watchdog_start() {
  static std::atomic<bool> started{false};
  const char* e = getenv("TDIST_SIM_WATCHDOG_S");
  if (e == nullptr || started.exchange(true)) return;
  static int secs = atoi(e);
  pthread_t t;
  pthread_create(&t, nullptr, [](void*) -> void* {
    timespec ts{secs, 0};
    nanosleep(&ts, nullptr);
    std::lock_guard<std::mutex> g(g_live_mu);
    for (CtaRun* r : g_live) {
      int done = 0, bar = 0, warp = 0, run = 0;
      for (auto& f : r->fibers) (f.done ? done : f.at_barrier ? bar : f.at_warp ? warp : run)++;
      fprintf(stderr, "sim watchdog: rank %d cta %d: %d done, %d at __syncthreads, %d at a warp barrier, %d runnable (current %d)\n",
              r->rank, r->cta, done, bar, warp, run, r->cur);
    }
    abort();
    return nullptr;
  }, nullptr);
  pthread_detach(t);
}

inline void fiber_main() {
  CtaRun* r = tl_cta;
  (*r->kernel)();
  r->fibers[r->cur].done = true;  // returning resumes the scheduler through uc_link
}

inline void* cta_thread(void* p) {
  CtaRun* r = static_cast<CtaRun*>(p);
  tl_cta = r;
  sim_cta = &r->state;
  sim_rank = r->rank;
  threadIdx.y = threadIdx.z = 0;  // (the dim3 default is 1, right for the extents only)
  blockDim.x = r->threads;
  if (r->gx > 0) {  // 3-D grid: x fastest
    blockIdx.x = r->cta % r->gx;
    blockIdx.y = (r->cta / r->gx) % r->gy;
    blockIdx.z = r->cta / (r->gx * r->gy);
    gridDim.x = r->gx;
    gridDim.y = r->gy;
    gridDim.z = r->gz;
  } else {
    blockIdx.x = r->cta;
    blockIdx.y = blockIdx.z = 0;
    gridDim.x = r->ctas;
    gridDim.y = gridDim.z = 1;
  }
  sim_rng = 2654435761u * unsigned(r->rank * 100003 + r->cta * 1009 + 1);
  r->fibers.resize(r->threads);
  {
    std::lock_guard<std::mutex> g(g_live_mu);
    g_live.push_back(r);
  }
  r->warp_scratch.assign(size_t(r->threads + 31) / 32 * 2048, 0);
  for (int t = 0; t < r->threads; ++t) {
    Fiber& f = r->fibers[t];
    f.stack.reset(new char[kFiberStack]);
    getcontext(&f.ctx);
    f.ctx.uc_stack.ss_sp = f.stack.get();
    f.ctx.uc_stack.ss_size = kFiberStack;
    f.ctx.uc_link = &r->sched;
    makecontext(&f.ctx, fiber_main, 0);
  }
  for (;;) {
    int alive = 0, waiting = 0;
    for (int t = 0; t < r->threads; ++t) {
      Fiber& f = r->fibers[t];
      if (f.done) continue;
      ++alive;
      if (f.at_barrier) {
        ++waiting;
        continue;
      }
      if (f.at_warp) continue;
      r->cur = t;
      threadIdx.x = t;
      swapcontext(&r->sched, &f.ctx);  // runs fiber t until it reaches a barrier or returns
      if (!f.done && f.at_barrier) ++waiting;
      if (f.done) --alive;
    }
    if (alive == 0) {
      std::lock_guard<std::mutex> g(g_live_mu);
      g_live.erase(std::find(g_live.begin(), g_live.end(), r));
      break;
    }
    if (waiting == alive)  // everybody has arrived: open the barrier
      for (auto& f : r->fibers) f.at_barrier = false;
    // warp barriers: a waiting lane is released when every live lane its mask names waits with the same mask (two
    // half-warps that shuffle with 0x0000ffff / 0xffff0000 meet independently of each other, as on the device)
    for (int w0 = 0; w0 < r->threads; w0 += 32) {
      const int w1 = std::min(w0 + 32, r->threads);
      for (int t = w0; t < w1; ++t) {
        const Fiber& f = r->fibers[t];
        if (f.done || !f.at_warp) continue;
        const unsigned m = f.warp_mask;
        bool open = true;
        for (int u = w0; u < w1 && open; ++u) {
          const Fiber& g = r->fibers[u];
          if (!((m >> (u - w0)) & 1u) || g.done) continue;
          open = g.at_warp && g.warp_mask == m;
        }
        if (open)
          for (int u = w0; u < w1; ++u)
            if ((m >> (u - w0)) & 1u) r->fibers[u].at_warp = false;
      }
    }
  }
  return nullptr;
}

}  // namespace sim

// called by the kernel sources through the CUDA stand-ins of stub/cuda_runtime.h
void sim_syncthreads() {
  sim::CtaRun* r = sim::tl_cta;
  sim::Fiber& f = r->fibers[r->cur];
  f.at_barrier = true;
  swapcontext(&f.ctx, &r->sched);
}
void sim_syncwarp(unsigned mask) {
  sim::CtaRun* r = sim::tl_cta;
  sim::Fiber& f = r->fibers[r->cur];
  f.warp_mask = mask;
  f.at_warp = true;
  swapcontext(&f.ctx, &r->sched);
}
void* sim_warp_scratch() {
  sim::CtaRun* r = sim::tl_cta;
  return r->warp_scratch.data() + size_t(r->cur / 32) * 2048;
}
void sim_yield() {  // a polling thread lets its siblings run: on the GPU they run in parallel
  sim::CtaRun* r = sim::tl_cta;
  if (r == nullptr) return;
  swapcontext(&r->fibers[r->cur].ctx, &r->sched);
}

namespace sim {

// One kernel launch of one rank: returns when every CTA has finished (launches of a rank are sequential, like a stream).
inline void launch(int rank, int ctas, int threads, const std::function<void()>& kernel, int gx = 0, int gy = 1, int gz = 1) {
  watchdog_start();
  std::vector<std::unique_ptr<CtaRun>> runs(ctas);
  std::vector<pthread_t> th(ctas);
  for (int c = 0; c < ctas; ++c) {
    runs[c].reset(new CtaRun);
    runs[c]->kernel = &kernel;
    runs[c]->rank = rank;
    runs[c]->cta = c;
    runs[c]->ctas = ctas;
    runs[c]->threads = threads;
    runs[c]->gx = gx;
    runs[c]->gy = gy;
    runs[c]->gz = gz;
    if (pthread_create(&th[c], nullptr, cta_thread, runs[c].get()) != 0) {
      perror("pthread_create");
      abort();
    }
  }
  for (auto& t : th) pthread_join(t, nullptr);
}

// The node: W heaps in one allocation, zero-filled once (flags start at zero and are never reset afterwards).
struct Node {
  int world;
  size_t bytes;
  char* heap;
  char* mc_window = nullptr;  // address range that stands for the NVLS multicast mapping (never dereferenced)
  Node(int w, size_t per_rank, bool multicast = false) : world(w), bytes((per_rank + 4095) / 4096 * 4096) {
    heap = static_cast<char*>(aligned_alloc(4096, bytes * w));
    std::memset(heap, 0, bytes * w);
    if (multicast && w > 1) {
      mc_window = static_cast<char*>(aligned_alloc(4096, bytes));
      sim_mc = SimMc{mc_window, heap, bytes, w};
    }
  }
  ~Node() {
    free(heap);
    free(mc_window);
  }
  char* base(int r) const { return heap + size_t(r) * bytes; }
  tdist::SymmCtx ctx(int r) const {
    tdist::SymmCtx c;
    for (int q = 0; q < TDIST_MAX_RANKS; ++q) c.base[q] = q < world ? base(q) : nullptr;
    c.mc_base = mc_window;  // nullptr: kernels take their point-to-point paths
    c.bytes = bytes;
    c.rank = r;
    c.world = world;
    return c;
  }
};

// 3-D grid of gx x gy x gz CTAs (single-GPU kernels such as the split-KV decode attention)
inline void launch3(int gx, int gy, int gz, int threads, const std::function<void()>& kernel) {
  launch(0, gx * gy * gz, threads, kernel, gx, gy, gz);
}

// Every rank runs `driver(rank)` on its own OS thread: ranks issue their launches independently of each other.
inline void run_ranks(int world, const std::function<void(int)>& driver) {
  struct Arg {
    const std::function<void(int)>* f;
    int r;
  };
  std::vector<Arg> args(world);
  std::vector<pthread_t> th(world);
  for (int r = 0; r < world; ++r) {
    args[r] = Arg{&driver, r};
    pthread_create(&th[r], nullptr, [](void* p) -> void* {
      auto* a = static_cast<Arg*>(p);
      (*a->f)(a->r);
      return nullptr;
    }, &args[r]);
  }
  for (auto& t : th) pthread_join(t, nullptr);
}

// deterministic pseudo-random test data in [-1, 1)
inline float unit_rand(unsigned& s) {
  s = s * 1664525u + 1013904223u;
  return float(int(s >> 8) & 0xffff) / 32768.f - 1.f;
}

}  // namespace sim
