// Host stand-in for csrc/kernels/common.cuh (+ device/symm.cuh, device/ptx.cuh) used by the protocol simulator
// (csrc/sim/protocol_sim.cc): the SAME kernel source is compiled for the CPU (one OS thread per CTA, its CUDA threads as
// fibers), with
//   * payload loads / stores (ld_v4, st_v4, ld_nc_v4)            -> plain, non-atomic accesses
//   * flags (notify, wait_*_sys, st_release_sys, ld_acquire_sys)  -> C++ atomics with the SAME memory orders
//   * fence()                                                     -> compiler barrier (the release / acquire pair carries the order)
//   * __syncthreads()                                             -> yield until all fibers of the CTA have arrived
// so that ThreadSanitizer checks what the protocol claims: every payload access is ordered by a release / acquire
// chain through the flags. It finds logic errors (a missing barrier or fence, a wrong sequence number, a slot reused too
// early, a deadlock); it does not model the GPU's weaker scopes or proxies.
#pragma once
#include <pthread.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include <math.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define TDIST_MAX_RANKS 16

struct int4 {
  int x, y, z, w;
};
inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
struct sim_dim3 {
  unsigned x = 1, y = 1, z = 1;
};
struct SimCta {
  uint32_t sh_phase = 0;  // phase_begin()'s shared word
  // __shared__ arrays of the kernel source: the staging step rewrites `__shared__ T name[..];` into a reference to a
  // per-CTA buffer obtained here (same pointer for every thread of the CTA)
  std::mutex mu;
  std::map<std::string, std::vector<char>> shared;
  void* shared_buf(size_t bytes, const char* key) {
    std::lock_guard<std::mutex> g(mu);
    auto& v = shared[key];
    if (v.size() < bytes) v.resize(bytes);
    return v.data();
  }
};
extern thread_local sim_dim3 threadIdx, blockIdx, blockDim, gridDim;
extern thread_local SimCta* sim_cta;
extern thread_local unsigned sim_rng;
void sim_syncthreads();  // csrc/sim/sim_runtime.h: yield this fiber until every CUDA thread of the CTA has arrived
inline void __syncthreads() { sim_syncthreads(); }
// schedule noise: a flag poll or a flag write is where interleavings matter
inline void sim_jitter() {
  sim_rng = sim_rng * 1664525u + 1013904223u;
  if ((sim_rng >> 28) == 0) sched_yield();
}
// TDIST_SIM_SLOW_RANK=r: every flag write of rank r is delayed by TDIST_SIM_SLOW_MS (default 2) milliseconds: a straggler
inline void sim_straggle(int rank) {
  static const int slow = getenv("TDIST_SIM_SLOW_RANK") ? atoi(getenv("TDIST_SIM_SLOW_RANK")) : -1;
  static const long ms = getenv("TDIST_SIM_SLOW_MS") ? atol(getenv("TDIST_SIM_SLOW_MS")) : 2;
  if (rank == slow) {
    timespec ts{ms / 1000, (ms % 1000) * 1000000};
    nanosleep(&ts, nullptr);
  }
}
template <typename T>
inline T min(T a, T b) { return a < b ? a : b; }
#define SIM_SHARED(type_t, name) type_t& name = *static_cast<type_t*>(sim_cta->shared_buf(sizeof(type_t), #name))

// bf16 with round-to-nearest-even, as the device conversions
struct __nv_bfloat16 {
  uint16_t v;
};
inline __nv_bfloat16 __float2bfloat16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return __nv_bfloat16{uint16_t((u >> 16) | 0x40)};  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return __nv_bfloat16{uint16_t(u >> 16)};
}
inline float __bfloat162float(__nv_bfloat16 b) {
  const uint32_t u = uint32_t(b.v) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

namespace tdist {

struct SymmCtx {
  char* base[TDIST_MAX_RANKS];
  char* mc_base;
  unsigned long long bytes;
  int rank;
  int world;
};
enum SignalOp : int { SIGNAL_SET = 9, SIGNAL_ADD = 10 };

template <typename T>
inline T* symm_at(const SymmCtx& c, T* local, int peer) {
  return reinterpret_cast<T*>(reinterpret_cast<uintptr_t>(c.base[peer]) +
                              (reinterpret_cast<uintptr_t>(local) - reinterpret_cast<uintptr_t>(c.base[c.rank])));
}

template <typename T>
inline T* mc_ptr(const SymmCtx& c, T* local) {
  return reinterpret_cast<T*>(reinterpret_cast<uintptr_t>(c.mc_base) +
                              (reinterpret_cast<uintptr_t>(local) - reinterpret_cast<uintptr_t>(c.base[c.rank])));
}

namespace ptx {
inline int4 ld_v4(const void* p) { return *static_cast<const int4*>(p); }
inline int4 ld_nc_v4(const void* p) { return *static_cast<const int4*>(p); }
inline void st_v4(void* p, const int4& v) { *static_cast<int4*>(p) = v; }
// the simulated heap has no multicast mapping (ctx.mc_base == nullptr): the kernels take their point-to-point paths
inline void multimem_st_v4(void*, const int4&) { abort(); }
inline int4 multimem_ld_reduce_bf16x8(const void*) { abort(); }
}  // namespace ptx

struct BF16 {};
template <typename Tag>
struct Pack;
template <>
struct Pack<BF16> {
  static void add(float (&acc)[8], const int4& v) {
    const uint16_t* h = reinterpret_cast<const uint16_t*>(&v);
    for (int i = 0; i < 8; ++i) acc[i] += __bfloat162float(__nv_bfloat16{h[i]});
  }
  static int4 cvt(const float (&acc)[8]) {
    int4 r;
    uint16_t* h = reinterpret_cast<uint16_t*>(&r);
    for (int i = 0; i < 8; ++i) h[i] = __float2bfloat16(acc[i]).v;
    return r;
  }
};

// per-context launch counter, as csrc/kernels/common.cuh: read by every CTA at the start of a launch, advanced by the last
// CTA to leave (launches of one rank are sequential, as on a stream)
struct PhaseState {
  uint32_t* phase;
  uint32_t* done;
};
inline uint32_t phase_begin(const PhaseState& s) {
  if (threadIdx.x == 0) sim_cta->sh_phase = __atomic_load_n(s.phase, __ATOMIC_ACQUIRE);
  __syncthreads();
  return sim_cta->sh_phase;
}
inline void phase_end(const PhaseState& s, uint32_t base, uint32_t steps) {
  __syncthreads();
  if (threadIdx.x == 0) {
    if (__atomic_fetch_add(s.done, 1u, __ATOMIC_ACQ_REL) == gridDim.x - 1) {
      __atomic_store_n(s.done, 0u, __ATOMIC_RELAXED);
      __atomic_store_n(s.phase, base + steps, __ATOMIC_RELEASE);
    }
  }
}

#ifdef TDIST_SIM_BREAK_RELEASE
// Negative control: the flag is written without release ordering and the fence does nothing. ThreadSanitizer must report the
// payload race this creates (the simulator's self-test).
inline void fence() {}
inline void notify(const SymmCtx& c, uint32_t* flag, int peer, uint32_t value, SignalOp = SIGNAL_SET) {
  sim_jitter();
  __atomic_store_n(symm_at(c, flag, peer), value, __ATOMIC_RELAXED);
}
inline void wait_ge_sys(const uint32_t* flag, uint32_t value) {
  while (static_cast<int32_t>(__atomic_load_n(flag, __ATOMIC_RELAXED) - value) < 0) sim_jitter();
}
#else
// ThreadSanitizer does not model stand-alone fences; the ordering it checks is the release store in notify() and the acquire
// load in wait_ge_sys(), which is also what the device primitives guarantee on their own (st.release.sys / ld.acquire.sys).
inline void fence() { std::atomic_signal_fence(std::memory_order_seq_cst); }
inline void notify(const SymmCtx& c, uint32_t* flag, int peer, uint32_t value, SignalOp op = SIGNAL_SET) {
  sim_jitter();
  sim_straggle(c.rank);
  if (op == SIGNAL_SET)
    __atomic_store_n(symm_at(c, flag, peer), value, __ATOMIC_RELEASE);
  else
    __atomic_fetch_add(symm_at(c, flag, peer), value, __ATOMIC_RELEASE);
}
inline void wait_ge_sys(const uint32_t* flag, uint32_t value) {
  const auto t0 = std::chrono::steady_clock::now();
  unsigned polls = 0;
  while (static_cast<int32_t>(__atomic_load_n(flag, __ATOMIC_ACQUIRE) - value) < 0) {
    sim_jitter();
    if ((++polls & 0xFFFFu) == 0 && std::chrono::steady_clock::now() - t0 > std::chrono::seconds(60)) {
      fprintf(stderr, "sim: wait timed out: flag %p holds %u, waiting for %u (rank-local block %u thread %u)\n", (const void*)flag,
              __atomic_load_n(flag, __ATOMIC_RELAXED), value, blockIdx.x, threadIdx.x);
      abort();
    }
  }
}
#endif

}  // namespace tdist
