// common.cuh — shared helpers for the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/tdist.h"
#include "../device/ptx.cuh"
#include "../device/symm.cuh"

namespace tdist {
void set_error(const std::string& msg);

static_assert(sizeof(SymmCtx) == sizeof(tdist_symm_ctx), "C and C++ views of the symmetric context must match");

inline SymmCtx to_ctx(const tdist_symm_ctx* c) {
  SymmCtx out;
  for (int i = 0; i < TDIST_MAX_RANKS; ++i) out.base[i] = c->base[i];
  out.mc_base = c->mc_base;
  out.bytes = c->bytes;
  out.rank = c->rank;
  out.world = c->world;
  return out;
}

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    return -1;
  }
  return 0;
}

// ---- 16-byte packed arithmetic in fp32, per dtype tag
struct BF16 {};
struct F16 {};
struct F32 {};

template <typename Tag>
struct Pack;  // add(acc[.], int4) accumulates; cvt(acc) packs back

template <>
struct Pack<BF16> {
  static constexpr int kElems = 8;
  __device__ static void add(float (&acc)[8], const int4& v) {
    const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = ptx::unpack_bf16x2(u[i]);
      acc[2 * i] += f.x;
      acc[2 * i + 1] += f.y;
    }
  }
  __device__ static int4 cvt(const float (&acc)[8]) {
    int4 r;
    r.x = ptx::pack_bf16x2(acc[0], acc[1]);
    r.y = ptx::pack_bf16x2(acc[2], acc[3]);
    r.z = ptx::pack_bf16x2(acc[4], acc[5]);
    r.w = ptx::pack_bf16x2(acc[6], acc[7]);
    return r;
  }
  __device__ static int4 mc_ld_reduce(const void* mc) { return ptx::multimem_ld_reduce_bf16x8(mc); }
};
template <>
struct Pack<F16> {
  static constexpr int kElems = 8;
  __device__ static void add(float (&acc)[8], const int4& v) {
    const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = ptx::unpack_f16x2(u[i]);
      acc[2 * i] += f.x;
      acc[2 * i + 1] += f.y;
    }
  }
  __device__ static int4 cvt(const float (&acc)[8]) {
    int4 r;
    r.x = ptx::pack_f16x2(acc[0], acc[1]);
    r.y = ptx::pack_f16x2(acc[2], acc[3]);
    r.z = ptx::pack_f16x2(acc[4], acc[5]);
    r.w = ptx::pack_f16x2(acc[6], acc[7]);
    return r;
  }
  __device__ static int4 mc_ld_reduce(const void* mc) { return ptx::multimem_ld_reduce_f16x8(mc); }
};
template <>
struct Pack<F32> {
  static constexpr int kElems = 8;  // acc array kept at 8 for a uniform signature; 4 used
  __device__ static void add(float (&acc)[8], const int4& v) {
    acc[0] += __int_as_float(v.x);
    acc[1] += __int_as_float(v.y);
    acc[2] += __int_as_float(v.z);
    acc[3] += __int_as_float(v.w);
  }
  __device__ static int4 cvt(const float (&acc)[8]) {
    int4 r;
    r.x = __float_as_int(acc[0]);
    r.y = __float_as_int(acc[1]);
    r.z = __float_as_int(acc[2]);
    r.w = __float_as_int(acc[3]);
    return r;
  }
  __device__ static int4 mc_ld_reduce(const void* mc) { return ptx::multimem_ld_reduce_f32x4(mc); }
};

// Launch-wide monotone phase: every CTA reads the same base value at entry,
// the last CTA to leave bumps it by `steps`. Keeps multi-CTA cross-GPU flag
// protocols reset-free and CUDA-graph capturable (no host-side counters).
struct PhaseState {
  uint32_t* phase;  // device counter (any memory)
  uint32_t* done;   // device counter, 0 between launches
};
__device__ __forceinline__ uint32_t phase_begin(const PhaseState& s) {
  __shared__ uint32_t sh_phase;
  if (threadIdx.x == 0) sh_phase = ptx::ld_volatile(s.phase);
  __syncthreads();
  return sh_phase;
}
__device__ __forceinline__ void phase_end(const PhaseState& s, uint32_t base, uint32_t steps) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t nb = gridDim.x * gridDim.y * gridDim.z;
    __threadfence();
    if (atomicAdd(s.done, 1u) == nb - 1) {
      *s.done = 0;
      *s.phase = base + steps;
      __threadfence();
    }
  }
}

}  // namespace tdist

#define TDIST_DISPATCH_DTYPE(dtype, ...)                         \
  switch (dtype) {                                               \
    case TDIST_BF16: { using Tag = ::tdist::BF16; __VA_ARGS__; break; } \
    case TDIST_F16: { using Tag = ::tdist::F16; __VA_ARGS__; break; }   \
    case TDIST_F32: { using Tag = ::tdist::F32; __VA_ARGS__; break; }   \
    default: ::tdist::set_error("unsupported dtype"); return -1; \
  }
