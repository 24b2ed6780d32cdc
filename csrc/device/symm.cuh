// symm.cuh — device-side distributed primitives over the symmetric heap.
//
// This header is the CUDA replacement for the reference's compiler-level
// primitives (language/distributed_ops.py:56-111; lowering semantics in
// lib/Conversion/TritonDistributedToLLVM/NVIDIA/DistributedOpToLLVM.cpp:179-416)
// and for the subset of the NVSHMEM device API its kernels use
// (language/extra/libshmem_device.py:28-342): rank / num_ranks / symm_at /
// notify / wait, putmem / getmem (block and warp scope), signal_op,
// signal_wait_until, fence / quiet, remote_ptr / remote_mc_ptr and a
// cross-GPU barrier.
//
// There is no proxy and no NIC path: every peer heap is mapped into this
// process' address space (VMM or cudaIpc), so a "put" is a store to a peer
// address and a "get" is a load from one. `consume_token` has no equivalent
// because ordering is explicit here: wait() ends in an acquire load, and the
// data loads that follow it in program order are ordered after it.
#pragma once
#include <stdio.h>
#include "ptx.cuh"

#define TDIST_MAX_RANKS 16

namespace tdist {

// Passed BY VALUE to every distributed kernel (fits easily in param space).
struct SymmCtx {
  char* base[TDIST_MAX_RANKS];  // base[r] = where rank r's heap is mapped in *this* process
  char* mc_base;                // NVLS multicast mapping of the heap (nullptr when unavailable)
  unsigned long long bytes;     // per-rank heap size
  int rank;
  int world;
};

enum SignalOp : int { SIGNAL_SET = 9, SIGNAL_ADD = 10 };  // numeric values follow NVSHMEM_SIGNAL_{SET,ADD}
enum CmpOp : int { CMP_EQ = 0, CMP_NE = 1, CMP_GT = 2, CMP_LE = 3, CMP_LT = 4, CMP_GE = 5 };

__device__ __forceinline__ int rank(const SymmCtx& c) { return c.rank; }
__device__ __forceinline__ int num_ranks(const SymmCtx& c) { return c.world; }

// Translate a pointer into my heap to the same object on `peer`.
template <typename T>
__device__ __forceinline__ T* symm_at(const SymmCtx& c, T* local, int peer) {
  return reinterpret_cast<T*>(reinterpret_cast<uintptr_t>(c.base[peer]) +
                              (reinterpret_cast<uintptr_t>(local) - reinterpret_cast<uintptr_t>(c.base[c.rank])));
}
template <typename T>
__device__ __forceinline__ T* remote_ptr(const SymmCtx& c, T* local, int peer) {
  return symm_at(c, local, peer);
}
// Multicast alias of a heap pointer (stores reach every rank, ld_reduce sums over ranks).
template <typename T>
__device__ __forceinline__ T* mc_ptr(const SymmCtx& c, T* local) {
  return reinterpret_cast<T*>(reinterpret_cast<uintptr_t>(c.mc_base) +
                              (reinterpret_cast<uintptr_t>(local) - reinterpret_cast<uintptr_t>(c.base[c.rank])));
}

// ---------------------------------------------------------------- notify / wait (32-bit flags)
// notify: release-ordered signal on `peer`'s copy of `flag`. Call from ONE
// thread after the CTA's payload stores are complete (bar.sync first).
__device__ __forceinline__ void notify(const SymmCtx& c, uint32_t* flag, int peer, uint32_t value,
                                       SignalOp op = SIGNAL_SET) {
  uint32_t* dst = symm_at(c, flag, peer);
  if (op == SIGNAL_SET)
    ptx::st_release_sys(dst, value);
  else
    ptx::red_add_release_sys(dst, value);
}
__device__ __forceinline__ void notify_local(uint32_t* flag, uint32_t value) { ptx::st_release_sys(flag, value); }

// Debug aid (the reference warns that a missed signal hangs forever, common_ops.py:58-60): building with
// -DTDIST_SPIN_LIMIT=<iterations> (python build.py --spin-limit N) turns an endless spin into a device printf naming
// the flag, the value seen and the value awaited, followed by a trap, so a protocol bug shows up as a CUDA error with a
// message instead of a hung job. Off by default: the check costs a counter in every polling loop.
#ifdef TDIST_SPIN_LIMIT
#define TDIST_SPIN_GUARD_DECL unsigned long long tdist_spins_ = 0
#define TDIST_SPIN_GUARD(flag, have, want)                                                                             \
  if (++tdist_spins_ > static_cast<unsigned long long>(TDIST_SPIN_LIMIT)) {                                            \
    printf("tdist: spin limit: flag %p holds %u, waiting for %u (block %d thread %d)\n", (const void*)(flag), (have),   \
           (want), int(blockIdx.x), int(threadIdx.x));                                                                 \
    __trap();                                                                                                          \
  }
#else
// Always-on bound (release builds): every 2^20 polls of one flag the wall clock is read; a wait that has lasted longer
// than TDIST_SPIN_TIMEOUT_S seconds (default 60; a healthy wait is micro- to milliseconds) prints the flag and traps, so
// a dead or wedged peer turns into a CUDA error within a minute instead of hanging the node until the NCCL watchdog
// (1800 s). Cost: one counter increment per poll.
#ifndef TDIST_SPIN_TIMEOUT_S
#define TDIST_SPIN_TIMEOUT_S 60
#endif
#define TDIST_SPIN_GUARD_DECL    \
  unsigned int tdist_spins_ = 0; \
  unsigned long long tdist_t0_ = 0
#define TDIST_SPIN_GUARD(flag, have, want)                                                                             \
  if ((++tdist_spins_ & 0xFFFFFu) == 0) {                                                                              \
    const unsigned long long now_ = ::tdist::ptx::globaltimer();                                                       \
    if (tdist_t0_ == 0) {                                                                                              \
      tdist_t0_ = now_;                                                                                                \
    } else if (now_ - tdist_t0_ > 1000000000ull * TDIST_SPIN_TIMEOUT_S) {                                              \
      printf("tdist: wait timed out after %d s: flag %p holds %u, waiting for %u (block %d thread %d)\n",              \
             int(TDIST_SPIN_TIMEOUT_S), (const void*)(flag), (have), (want), int(blockIdx.x), int(threadIdx.x));       \
      __trap();                                                                                                        \
    }                                                                                                                  \
  }
#endif

// Spin until *flag == value (acquire, system scope). One lane polls.
__device__ __forceinline__ void wait_eq_sys(const uint32_t* flag, uint32_t value) {
  TDIST_SPIN_GUARD_DECL;
  uint32_t v;
  while ((v = ptx::ld_acquire_sys(flag)) != value) {
    TDIST_SPIN_GUARD(flag, v, value)
  }
}
__device__ __forceinline__ void wait_ge_sys(const uint32_t* flag, uint32_t value) {
  TDIST_SPIN_GUARD_DECL;
  uint32_t v;
  while (static_cast<int32_t>((v = ptx::ld_acquire_sys(flag)) - value) < 0) {
    TDIST_SPIN_GUARD(flag, v, value)
  }
}
// Same wait, polling with RELAXED loads and ending with one acquire load of the flag. Inside a kernel that keeps many
// loads in flight (TMA pipelines, NVLS loads) an acquire in every poll iteration is expensive: a system-scope acquire
// also waits for the SM's other outstanding memory traffic (measured on the fused GEMM+RS at 8 ranks, profiles/README.md).
__device__ __forceinline__ void wait_ge_sys_poll_relaxed(const uint32_t* flag, uint32_t value) {
  TDIST_SPIN_GUARD_DECL;
  uint32_t v;
  while (static_cast<int32_t>((v = ptx::ld_relaxed_sys(flag)) - value) < 0) {
    TDIST_SPIN_GUARD(flag, v, value)
  }
  (void)ptx::ld_acquire_sys(flag);
}
// Warp-collective wait on n consecutive flags (n <= 32): the first n lanes of
// the warp each poll one flag, then the warp re-converges — this is exactly
// how the reference lowers dl.wait (DistributedOpToLLVM.cpp:179-206).
__device__ __forceinline__ void wait_warp(const uint32_t* flags, int n, uint32_t value) {
  const int lane = ptx::laneid();
  if (lane < n) wait_eq_sys(flags + lane, value);
  __syncwarp();
}
__device__ __forceinline__ void signal_wait_until(const uint32_t* flag, CmpOp cmp, uint32_t value) {
  for (;;) {
    const uint32_t v = ptx::ld_acquire_sys(flag);
    bool ok;
    switch (cmp) {
      case CMP_EQ: ok = v == value; break;
      case CMP_NE: ok = v != value; break;
      case CMP_GT: ok = v > value; break;
      case CMP_LE: ok = v <= value; break;
      case CMP_LT: ok = v < value; break;
      default: ok = v >= value; break;
    }
    if (ok) return;
  }
}
__device__ __forceinline__ void fence() { ptx::fence_acq_rel_sys(); }
__device__ __forceinline__ void quiet() { ptx::fence_sc_sys(); }

// ---------------------------------------------------------------- put / get
// CTA-cooperative copy of `bytes` (multiple of 16, both sides 16 B aligned).
// dst may be a peer-mapped address (put) or src may be one (get).
template <int kUnroll = 4>
__device__ __forceinline__ void copy_block(void* dst, const void* src, size_t bytes) {
  const size_t n = bytes >> 4;
  int4* d = reinterpret_cast<int4*>(dst);
  const int4* s = reinterpret_cast<const int4*>(src);
  size_t i = threadIdx.x;
  const size_t stride = blockDim.x;
  for (; i + (kUnroll - 1) * stride < n; i += kUnroll * stride) {
    int4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) v[u] = ptx::ld_v4(s + i + u * stride);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) ptx::st_v4(d + i + u * stride, v[u]);
  }
  for (; i < n; i += stride) ptx::st_v4(d + i, ptx::ld_v4(s + i));
}
template <int kUnroll = 4>
__device__ __forceinline__ void copy_warp(void* dst, const void* src, size_t bytes) {
  const size_t n = bytes >> 4;
  int4* d = reinterpret_cast<int4*>(dst);
  const int4* s = reinterpret_cast<const int4*>(src);
  size_t i = ptx::laneid();
  for (; i + (kUnroll - 1) * 32 < n; i += kUnroll * 32) {
    int4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) v[u] = ptx::ld_v4(s + i + u * 32);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) ptx::st_v4(d + i + u * 32, v[u]);
  }
  for (; i < n; i += 32) ptx::st_v4(d + i, ptx::ld_v4(s + i));
}
template <typename T>
__device__ __forceinline__ void putmem_block(const SymmCtx& c, T* dst_local, const T* src, size_t bytes, int peer) {
  copy_block(symm_at(c, dst_local, peer), src, bytes);
}
template <typename T>
__device__ __forceinline__ void getmem_block(const SymmCtx& c, T* dst, T* src_local, size_t bytes, int peer) {
  copy_block(dst, symm_at(c, src_local, peer), bytes);
}
template <typename T>
__device__ __forceinline__ void putmem_warp(const SymmCtx& c, T* dst_local, const T* src, size_t bytes, int peer) {
  copy_warp(symm_at(c, dst_local, peer), src, bytes);
}
// put + signal: payload by the whole CTA, then one thread publishes the flag.
template <typename T>
__device__ __forceinline__ void putmem_signal_block(const SymmCtx& c, T* dst_local, const T* src, size_t bytes,
                                                    uint32_t* flag, uint32_t value, SignalOp op, int peer) {
  copy_block(symm_at(c, dst_local, peer), src, bytes);
  __syncthreads();
  if (threadIdx.x == 0) {
    ptx::fence_acq_rel_sys();
    notify(c, flag, peer, value, op);
  }
}

// The rest of the NVSHMEM device vocabulary the reference exposes (language/extra/libshmem_device.py:28-342), as thin
// forms over the mapped peer pointers. There is no proxy thread and no non-blocking queue: a "_nbi" put is the same
// store sequence, completion = the fence / quiet that precedes the signal.
__device__ __forceinline__ int my_pe(const SymmCtx& c) { return c.rank; }
__device__ __forceinline__ int n_pes(const SymmCtx& c) { return c.world; }
__device__ __forceinline__ void int_p(const SymmCtx& c, int* dst_local, int value, int peer) {
  *symm_at(c, dst_local, peer) = value;
}
__device__ __forceinline__ void signal_op(const SymmCtx& c, uint32_t* flag, uint32_t value, SignalOp op, int peer) {
  notify(c, flag, peer, value, op);
}
template <typename T>
__device__ __forceinline__ void getmem_warp(const SymmCtx& c, T* dst, T* src_local, size_t bytes, int peer) {
  copy_warp(dst, symm_at(c, src_local, peer), bytes);
}
template <typename T>
__device__ __forceinline__ void putmem_nbi_block(const SymmCtx& c, T* dst_local, const T* src, size_t bytes, int peer) {
  putmem_block(c, dst_local, src, bytes, peer);
}
template <typename T>
__device__ __forceinline__ void putmem_nbi_warp(const SymmCtx& c, T* dst_local, const T* src, size_t bytes, int peer) {
  putmem_warp(c, dst_local, src, bytes, peer);
}
// put + signal by one warp (the flag is published by lane 0 after the warp's stores)
template <typename T>
__device__ __forceinline__ void putmem_signal_warp(const SymmCtx& c, T* dst_local, const T* src, size_t bytes,
                                                   uint32_t* flag, uint32_t value, SignalOp op, int peer) {
  copy_warp(symm_at(c, dst_local, peer), src, bytes);
  __syncwarp();
  if (ptx::laneid() == 0) {
    ptx::fence_acq_rel_sys();
    notify(c, flag, peer, value, op);
  }
}
// broadcast: the root CTA writes `bytes` of its buffer into every peer's copy (one NVLS store stream when a multicast
// mapping exists, W - 1 unicast puts otherwise). Callers synchronise with a barrier or a signal as with NVSHMEM.
template <typename T>
__device__ __forceinline__ void broadcast_block(const SymmCtx& c, T* buf_local, size_t bytes, int root) {
  if (c.rank != root) return;
  if (c.mc_base != nullptr) {
    const int4* s = reinterpret_cast<const int4*>(buf_local);
    char* d = reinterpret_cast<char*>(mc_ptr(c, buf_local));
    for (size_t i = threadIdx.x; i < bytes / 16; i += blockDim.x) ptx::multimem_st_v4(d + i * 16, ptx::ld_v4(s + i));
  } else {
    for (int p = 1; p < c.world; ++p) putmem_block(c, buf_local, buf_local, bytes, (c.rank + p) % c.world);
  }
}
// fcollect (all-gather of equal blocks): every rank's CTA puts its `bytes` block at offset rank * bytes of every peer's
// destination buffer.
template <typename T>
__device__ __forceinline__ void fcollect_block(const SymmCtx& c, T* dst_local, const T* src, size_t bytes) {
  char* slot = reinterpret_cast<char*>(dst_local) + size_t(c.rank) * bytes;
  for (int p = 0; p < c.world; ++p) copy_block(symm_at(c, slot, (c.rank + p) % c.world), src, bytes);
}
// teams are strided subsets of the world (triton_dist/runtime/team.py): rank translation is arithmetic
__device__ __forceinline__ int team_translate_pe(int src_start, int src_stride, int pe_in_src, int dst_start, int dst_stride,
                                                 int dst_size) {
  const int w = src_start + pe_in_src * src_stride, d = w - dst_start;
  return (d < 0 || d % dst_stride != 0 || d / dst_stride >= dst_size) ? -1 : d / dst_stride;
}

// ---------------------------------------------------------------- cross-GPU barrier (1 CTA, >= world threads)
// Monotone-phase flag barrier: flags[src] on every rank holds the last phase
// in which `src` arrived. Caller passes a phase value that increases by one
// per barrier; nothing is ever reset, so there is no ABA window and the
// barrier is CUDA-graph capturable (phase lives in device memory, see
// barrier_all_kernel). Equivalent role: barrier_all_intra_node_* in the
// reference's common_ops.py:141-209.
__device__ __forceinline__ void barrier_all_block(const SymmCtx& c, uint32_t* flags, uint32_t phase) {
  __syncthreads();
  const int t = threadIdx.x;
  if (t < c.world) {
    ptx::fence_acq_rel_sys();
    ptx::st_release_sys(symm_at(c, flags + c.rank, t), phase);  // tell rank t that I arrived
    wait_ge_sys(flags + t, phase);  // wait until rank t arrived here
  }
  __syncthreads();
}

// ---------------------------------------------------------------- grid barrier (all CTAs co-resident)
// counter must start at 0 and is left at 0 mod 2^31 flip; same sign-bit trick
// as the reference's barrier_on_this_grid (common_ops.py:56-138) but with
// explicit release/acquire.
__device__ __forceinline__ void grid_barrier(uint32_t* counter) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t nb = gridDim.x * gridDim.y * gridDim.z;
    const bool master = (blockIdx.x + blockIdx.y + blockIdx.z) == 0;
    const uint32_t inc = master ? 0x80000000u - (nb - 1) : 1u;
    __threadfence();
    const uint32_t old = ptx::atom_add_acq_rel_gpu(counter, inc);
    while (((old ^ ptx::ld_acquire_gpu(counter)) & 0x80000000u) == 0) {
    }
    __threadfence();
  }
  __syncthreads();
}

}  // namespace tdist

// ---------------------------------------------------------------- small device utilities
namespace tdist {
// Binary search helpers (reference common_ops.py:256-344 bisect_left / bisect_right): first index in the sorted
// array a[0..n) whose element is >= v (left) or > v (right).
template <typename T>
__device__ __forceinline__ int bisect_left(const T* a, int n, T v) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (a[mid] < v)
      lo = mid + 1;
    else
      hi = mid;
  }
  return lo;
}
template <typename T>
__device__ __forceinline__ int bisect_right(const T* a, int n, T v) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (a[mid] <= v)
      lo = mid + 1;
    else
      hi = mid;
  }
  return lo;
}
}  // namespace tdist
