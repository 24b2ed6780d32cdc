// Host stand-in for <cuda_runtime.h> (and the CUDA language) used by the protocol simulator (csrc/sim): it lets the REAL
// csrc/kernels/common.cuh, csrc/device/symm.cuh and the kernel sources compile with g++. One OS thread runs one CTA, the
// CUDA threads of the CTA are fibers on it (sim_runtime.h); everything here is either a keyword that disappears, a type,
// or a hook into that scheduler.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

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
#define __align__(n) alignas(n)
#define __grid_constant__

struct int4 {
  int x, y, z, w;
};
struct uint2 {
  unsigned x, y;
};
struct float2 {
  float x, y;
};
struct alignas(16) float4 {
  float x, y, z, w;
};
inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }

struct sim_dim3 {
  unsigned x = 1, y = 1, z = 1;
};
// per-CTA state: __shared__ variables of the kernel sources (the staging step rewrites `__shared__ T name[..];` into a
// reference to a buffer obtained here: the same pointer for every thread of the CTA)
struct SimCta {
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
#define SIM_SHARED(type_t, name) type_t& name = *static_cast<type_t*>(sim_cta->shared_buf(sizeof(type_t), #name))
// `extern __shared__ T name[];` (dynamic shared memory): the 227 KB a CTA can opt into, 1024-byte aligned (TMA swizzle
// patterns are functions of the shared-memory address)
#define SIM_DYN_SHARED(type_t, name) \
  type_t* name = reinterpret_cast<type_t*>((reinterpret_cast<uintptr_t>(sim_cta->shared_buf(232448 + 1024, "dynamic")) + 1023) & ~uintptr_t(1023))

// scheduler hooks (csrc/sim/sim_runtime.h)
void sim_syncthreads();  // yield until every CUDA thread of the CTA has arrived
void sim_syncwarp(unsigned mask = 0xffffffffu);  // ... every lane of this thread's warp that `mask` names
void sim_yield();        // let the sibling threads of the CTA run (called from every flag poll)
inline void __syncthreads() { sim_syncthreads(); }
inline void __syncwarp(unsigned mask = 0xffffffffu) { sim_syncwarp(mask); }
void* sim_warp_scratch();  // 32 x 16 bytes shared by the lanes of the calling thread's warp
// warp shuffles: every lane publishes its value, the warp meets, every lane reads the source lane's value
template <typename T>
inline T __shfl_sync(unsigned mask, T v, int src_lane) {
  static_assert(sizeof(T) <= 16, "shuffle payload");
  char* w = static_cast<char*>(sim_warp_scratch());
  memcpy(w + (threadIdx.x & 31u) * 16, &v, sizeof(T));
  sim_syncwarp(mask);
  T r;
  memcpy(&r, w + (unsigned(src_lane) & 31u) * 16, sizeof(T));
  sim_syncwarp(mask);
  return r;
}
template <typename T>
inline T __shfl_xor_sync(unsigned m, T v, int lane_mask) { return __shfl_sync(m, v, int(threadIdx.x & 31u) ^ lane_mask); }
inline void __threadfence() { __atomic_signal_fence(__ATOMIC_SEQ_CST); }  // ordering is carried by the atomics around it
inline void __threadfence_block() { __atomic_signal_fence(__ATOMIC_SEQ_CST); }
inline void __threadfence_system() { __atomic_signal_fence(__ATOMIC_SEQ_CST); }
inline void __trap() { abort(); }

// schedule noise and stragglers
inline void sim_jitter() {
  sim_rng = sim_rng * 1664525u + 1013904223u;
  if ((sim_rng >> 28) == 0) sched_yield();
}
// TDIST_SIM_SLOW_RANK=r: every flag write of rank r is delayed by TDIST_SIM_SLOW_MS (default 2) milliseconds
inline void sim_straggle(int rank) {
  static const int slow = getenv("TDIST_SIM_SLOW_RANK") ? atoi(getenv("TDIST_SIM_SLOW_RANK")) : -1;
  static const long ms = getenv("TDIST_SIM_SLOW_MS") ? atol(getenv("TDIST_SIM_SLOW_MS")) : 2;
  if (rank == slow) {
    timespec ts{ms / 1000, (ms % 1000) * 1000000};
    nanosleep(&ts, nullptr);
  }
}
extern thread_local int sim_rank;  // rank of the launch this OS thread belongs to

template <typename T>
inline T min(T a, T b) { return a < b ? a : b; }
template <typename T>
inline T max(T a, T b) { return a > b ? a : b; }
inline size_t min(size_t a, unsigned long long b) { return a < b ? a : size_t(b); }

inline float __int_as_float(int v) {
  float f;
  memcpy(&f, &v, 4);
  return f;
}
inline int __float_as_int(float f) {
  int v;
  memcpy(&v, &f, 4);
  return v;
}
// device atomics: sequentially consistent here (on the device they are relaxed and fenced by the surrounding __threadfence)
inline unsigned atomicAdd(unsigned* p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline int atomicAdd(int* p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }

// bf16 / fp16 storage types with round-to-nearest-even conversions (what the device intrinsics do)
struct __nv_bfloat16 {
  uint16_t v;
};
inline __nv_bfloat16 __float2bfloat16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return __nv_bfloat16{uint16_t((u >> 16) | 0x40)};
  u += 0x7fffu + ((u >> 16) & 1u);
  return __nv_bfloat16{uint16_t(u >> 16)};
}
inline float __bfloat162float(__nv_bfloat16 b) {
  const uint32_t u = uint32_t(b.v) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// fp16 storage type (conversions: csrc/sim/shim/device/ptx.cuh has the bit-exact routines)
struct __half {
  uint16_t v;
};
uint16_t sim_f32_to_f16(float f);
float sim_f16_to_f32(uint16_t h);
inline __half __float2half(float f) { return __half{sim_f32_to_f16(f)}; }
inline float __half2float(__half h) { return sim_f16_to_f32(h.v); }
inline float rsqrtf(float x) { return 1.f / sqrtf(x); }
inline float __expf(float x) { return expf(x); }
inline float __frcp_rn(float x) { return 1.f / x; }
template <typename T>
inline T __ldcg(const T* p) { return *p; }
inline long long clock64() { return 0; }
inline float __logf(float x) { return logf(x); }
#define CUDART_INF_F (__builtin_inff())

// runtime API surface the launch helpers of common.cuh touch (never reached: the simulator calls the kernels directly)
typedef int cudaError_t;
typedef void* cudaStream_t;
constexpr cudaError_t cudaSuccess = 0;
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
// "device" memory of host code that is compiled into the simulator as it is (csrc/mega/builder.cc): host memory
enum cudaMemcpyKind { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16 };
template <typename T>
inline cudaError_t cudaMalloc(T** p, size_t n) {
  *p = static_cast<T*>(aligned_alloc(256, (n + 255) / 256 * 256));
  return *p != nullptr ? cudaSuccess : 2;
}
inline cudaError_t cudaFree(void* p) {
  free(p);
  return cudaSuccess;
}
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) {
  memcpy(d, s, n);
  return cudaSuccess;
}
inline cudaError_t cudaMemset(void* d, int v, size_t n) {
  memset(d, v, n);
  return cudaSuccess;
}
inline cudaError_t cudaGetDevice(int* d) {
  *d = 0;
  return cudaSuccess;
}
inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr, int) {
  *v = 148;
  return cudaSuccess;
}
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline const char* cudaGetErrorString(cudaError_t) { return "simulated"; }
