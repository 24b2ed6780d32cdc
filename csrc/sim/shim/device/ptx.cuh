// Host stand-in for csrc/device/ptx.cuh, used by the protocol simulator: the same function names, implemented with C++
// atomics and plain accesses so that ThreadSanitizer checks the protocols of the code that calls them.
//   payload:  ld_v4 / ld_nc_v4 / st_v4                      plain 16-byte accesses (what the sanitizer watches)
//   flags:    st_release_sys / red_add_release_sys          atomic, release        (relaxed with -DTDIST_SIM_BREAK_RELEASE:
//             ld_acquire_sys / ld_acquire_gpu               atomic, acquire         the negative control)
//             ld_relaxed_sys / ld_volatile                  atomic, relaxed
//   packets:  st_relaxed_sys_v4 / ld_volatile_v4            two relaxed 8-byte atomics {data, flag}: the LL protocol's unit
//   copies:   bulk and tiled TMA copies are performed at issue (mbarrier = count of completed phases)
//   fences:   compiler barriers — ThreadSanitizer does not model stand-alone fences; the order it checks is carried by the
//             release store that follows a fence and the acquire load that precedes the data loads, as on the device
//   NVLS:     multimem_st_* writes every rank's copy, multimem_ld_reduce_* reads every rank's copy and adds in fp32
// Every flag poll yields to the sibling CUDA threads of the CTA (they are fibers of one OS thread).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

struct SimMc {  // the simulated node's multicast window (set by sim::Node when multicast is enabled)
  char* mc_base = nullptr;
  char* heap = nullptr;
  size_t bytes = 0;
  int world = 0;
};
extern SimMc sim_mc;

namespace tdist {
namespace ptx {

#ifdef TDIST_SIM_BREAK_RELEASE
constexpr int kRelease = __ATOMIC_RELAXED, kAcquire = __ATOMIC_RELAXED;
#else
constexpr int kRelease = __ATOMIC_RELEASE, kAcquire = __ATOMIC_ACQUIRE;
#endif

inline int laneid() { return int(threadIdx.x & 31u); }
inline unsigned long long globaltimer() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (unsigned long long)ts.tv_sec * 1000000000ull + ts.tv_nsec;
}

// ---- payload
inline int4 ld_v4(const void* p) { return *static_cast<const int4*>(p); }
inline int4 ld_nc_v4(const void* p) { return *static_cast<const int4*>(p); }
inline void st_v4(void* p, const int4& v) { *static_cast<int4*>(p) = v; }

// ---- flags
inline void st_release_sys(uint32_t* p, uint32_t v) {
  sim_jitter();
  sim_straggle(sim_rank);
  __atomic_store_n(p, v, kRelease);
}
inline void st_release_gpu(uint32_t* p, uint32_t v) { st_release_sys(p, v); }
// A relaxed system-scope flag store is always the second half of "fence; store" in the sources (the fence — GPU or system
// scope — is what orders the payload before it). The sanitizer does not model stand-alone fences, so the pair is modelled
// as one release store.
inline void st_relaxed_sys(uint32_t* p, uint32_t v) {
  sim_jitter();
  sim_straggle(sim_rank);
  __atomic_store_n(p, v, kRelease);
}
inline void red_add_release_sys(uint32_t* p, uint32_t v) {
  sim_jitter();
  sim_straggle(sim_rank);
  __atomic_fetch_add(p, v, kRelease);
}
// "fence.acq_rel.gpu; red.relaxed.gpu" in the sources (the scoreboard signal of the megakernel): one release RMW, as above
inline void red_add_relaxed_gpu(uint32_t* p, uint32_t v) {
  sim_jitter();
  __atomic_fetch_add(p, v, kRelease);
}
inline uint32_t atom_add_acq_rel_gpu(uint32_t* p, uint32_t v) { return __atomic_fetch_add(p, v, __ATOMIC_ACQ_REL); }
inline uint32_t ld_acquire_sys(const uint32_t* p) {
  sim_yield();
  return __atomic_load_n(p, kAcquire);
}
inline uint32_t ld_acquire_gpu(const uint32_t* p) { return ld_acquire_sys(p); }
inline uint32_t ld_relaxed_sys(const uint32_t* p) {
  sim_yield();
  return __atomic_load_n(p, __ATOMIC_RELAXED);
}
inline uint32_t ld_volatile(const uint32_t* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }

// ---- fences
inline void fence_acq_rel_sys() { __atomic_signal_fence(__ATOMIC_SEQ_CST); }
inline void fence_acq_rel_gpu() { __atomic_signal_fence(__ATOMIC_SEQ_CST); }
inline void fence_sc_sys() { __atomic_signal_fence(__ATOMIC_SEQ_CST); }

// ---- 1-D bulk copies (the ring of the point-to-point AG push: global -> shared -> peer global, driven by one thread):
// performed at issue, like the tiled loads below; the store to the peer is a plain write, which is what the sanitizer checks
// against the consumer's reads
inline void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  memcpy(smem_dst, gsrc, bytes);
  ++*bar;
}
inline void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) { memcpy(gdst, smem_src, bytes); }
inline void bulk_commit() {}
template <int N>
inline void bulk_wait() {}
template <int N>
inline void bulk_wait_read() {}

// ---- mbarrier + tiled TMA loads, as far as the decode megakernel uses them: one arrival per phase, the transaction bytes
// of the phase delivered by ONE tensor load. The load is performed at issue (the copy engine of the simulator is infinitely
// fast), which completes the phase; the barrier word counts completed phases. All parties are fibers of one CTA.
inline uint32_t smem_u32(const void* p) { return uint32_t(reinterpret_cast<uintptr_t>(p)); }
inline void mbar_init(uint64_t* bar, uint32_t) { *bar = 0; }
inline void fence_barrier_init() {}
inline void fence_proxy_async() {}
inline void fence_proxy_async_shared() {}
inline void mbar_arrive_expect_tx(uint64_t*, uint32_t) {}
inline void mbar_wait(uint64_t* bar, uint32_t parity) {
  while ((*bar & 1u) == parity) sim_yield();  // the phase with this parity has not completed yet
}
inline void prefetch_tensormap(const void*) {}
inline void tma_prefetch_3d(const void*, int, int, int) {}
inline void prefetch_l2_bulk(const void*, uint32_t) {}
// box {box[0], box[1], box[2]} at element coordinates (c0, c1, c2); lands densely as [z][y][x]; elements outside the tensor
// are zero; with the 128 B swizzle the 16-byte chunks of every 128 B line are XORed with bits 7-9 of the line's address
inline void tma_load_3d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2) {
  SimTmap t;
  memcpy(&t, tmap, sizeof(t));
  char* dst = static_cast<char*>(smem_dst);
  const uint32_t row_bytes = t.box[0] * t.elt_bytes;
  if (t.swizzle == uint32_t(CU_TENSOR_MAP_SWIZZLE_128B) && (row_bytes != 128 || (reinterpret_cast<uintptr_t>(dst) & 127u) != 0)) {
    fprintf(stderr, "sim: tma_load_3d: a 128 B-swizzled box needs 128-byte rows and a 128-byte aligned destination\n");
    abort();
  }
  for (uint32_t z = 0; z < t.box[2]; ++z)
    for (uint32_t y = 0; y < t.box[1]; ++y) {
      char* line = dst + (size_t(z) * t.box[1] + y) * row_bytes;
      const uint32_t phase = t.swizzle == uint32_t(CU_TENSOR_MAP_SWIZZLE_128B) ? uint32_t(reinterpret_cast<uintptr_t>(line) >> 7) & 7u : 0u;
      const uint64_t gy = uint64_t(c1) + y, gz = uint64_t(c2) + z;
      const bool row_ok = c1 >= 0 && c2 >= 0 && gy < t.dims[1] && gz < t.dims[2];
      for (uint32_t x = 0; x < t.box[0]; ++x) {
        const uint64_t gx = uint64_t(c0) + x;
        uint32_t byte = x * t.elt_bytes;
        byte = (((byte >> 4) ^ phase) << 4) | (byte & 15u);
        if (row_ok && c0 >= 0 && gx < t.dims[0])
          memcpy(line + byte, t.base + gz * t.strides[1] + gy * t.strides[0] + gx * t.elt_bytes, t.elt_bytes);
        else
          memset(line + byte, 0, t.elt_bytes);
      }
    }
  ++*bar;
}

// ---- warp-level matrix instructions of the megakernel's linear task (the staging step replaces the two inline-asm
// statements of csrc/mega/megakernel.cu by calls to these): lanes exchange their operands through the warp scratch
// ldmatrix.sync.aligned.m8n8.x4.b16: lane l supplies the address of row l % 8 of matrix l / 8; register j of lane l receives
// elements 2 (l % 4), 2 (l % 4) + 1 of row l / 4 of matrix j
inline void sim_ldmatrix_x4(uint32_t (&a)[4], const void* smem_ptr) {
  char* w = static_cast<char*>(sim_warp_scratch()) + 512;
  const unsigned lane = threadIdx.x & 31u;
  memcpy(w + lane * 8, &smem_ptr, sizeof(void*));
  sim_syncwarp();
  for (unsigned j = 0; j < 4; ++j) {
    const char* row;
    memcpy(&row, w + (j * 8 + lane / 4) * 8, sizeof(void*));
    memcpy(&a[j], row + (lane % 4) * 4, 4);
  }
  sim_syncwarp();
}
// mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32: D[16 x 8] += A[16 x 16] B[16 x 8]. With g = lane / 4, t = lane % 4:
// a0 = A[g][2t, 2t+1], a1 = A[g+8][2t, 2t+1], a2 = A[g][2t+8, 2t+9], a3 = A[g+8][2t+8, 2t+9]; b0 = B[2t, 2t+1][g],
// b1 = B[2t+8, 2t+9][g]; d0, d1 = D[g][2t, 2t+1], d2, d3 = D[g+8][2t, 2t+1]
inline void sim_mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  char* w = static_cast<char*>(sim_warp_scratch()) + 512;
  const unsigned lane = threadIdx.x & 31u, g = lane / 4, t = lane % 4;
  uint32_t mine[6] = {a[0], a[1], a[2], a[3], b0, b1};
  memcpy(w + lane * 32, mine, sizeof(mine));
  sim_syncwarp();
  auto elem = [&](unsigned src_lane, unsigned reg, unsigned e) {
    uint32_t v;
    memcpy(&v, w + src_lane * 32 + reg * 4, 4);
    return __bfloat162float(__nv_bfloat16{uint16_t(e ? v >> 16 : v & 0xffffu)});
  };
  for (unsigned r = 0; r < 2; ++r)      // rows g, g + 8
    for (unsigned c = 0; c < 2; ++c) {  // columns 2t, 2t + 1
      float acc = d[r * 2 + c];
      const unsigned n = 2 * t + c;
      for (unsigned k = 0; k < 16; ++k) {
        const unsigned tk = (k % 8) / 2, e = k % 2, hi = k / 8;
        acc += elem(g * 4 + tk, r + 2 * hi, e) * elem(n * 4 + tk, 4 + hi, e);
      }
      d[r * 2 + c] = acc;
    }
  sim_syncwarp();
}

// ---- LL packets: a 16-byte packet is {d0, flag, d1, flag}; each 8-byte half carries its own flag
inline void st_relaxed_sys_v4(void* p, const int4& v) {
  uint64_t h[2];
  memcpy(h, &v, 16);
  __atomic_store_n(static_cast<uint64_t*>(p), h[0], __ATOMIC_RELAXED);
  __atomic_store_n(static_cast<uint64_t*>(p) + 1, h[1], __ATOMIC_RELAXED);
}
inline int4 ld_volatile_v4(const void* p) {
  sim_yield();
  uint64_t h[2];
  h[0] = __atomic_load_n(static_cast<const uint64_t*>(p), __ATOMIC_RELAXED);
  h[1] = __atomic_load_n(static_cast<const uint64_t*>(p) + 1, __ATOMIC_RELAXED);
  int4 v;
  memcpy(&v, h, 16);
  return v;
}

// ---- 16-bit float packing
inline uint16_t f32_to_f16(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  int e = int((x >> 23) & 0xff) - 127 + 15;
  uint32_t m = x & 0x7fffffu;
  if (e >= 31) return uint16_t(sign | 0x7c00u | (((x & 0x7fffffffu) > 0x7f800000u) ? 0x200u : 0u));
  if (e <= 0) {
    if (e < -10) return uint16_t(sign);
    m |= 0x800000u;
    const int shift = 14 - e;
    uint32_t r = m >> shift;
    const uint32_t rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (r & 1))) ++r;
    return uint16_t(sign | r);
  }
  uint32_t r = (uint32_t(e) << 10) | (m >> 13);
  const uint32_t rem = m & 0x1fffu;
  if (rem > 0x1000u || (rem == 0x1000u && (r & 1))) ++r;
  return uint16_t(sign | r);
}
inline float f16_to_f32(uint16_t h) {
  const uint32_t sign = uint32_t(h & 0x8000u) << 16;
  int e = (h >> 10) & 0x1f;
  uint32_t m = h & 0x3ffu, x;
  if (e == 0) {
    if (m == 0) {
      x = sign;
    } else {
      e = 1;
      while (!(m & 0x400u)) {
        m <<= 1;
        --e;
      }
      x = sign | (uint32_t(e + 112) << 23) | ((m & 0x3ffu) << 13);
    }
  } else if (e == 31) {
    x = sign | 0x7f800000u | (m << 13);
  } else {
    x = sign | (uint32_t(e + 112) << 23) | (m << 13);
  }
  float f;
  memcpy(&f, &x, 4);
  return f;
}
inline uint32_t pack_bf16x2(float lo, float hi) { return uint32_t(__float2bfloat16(lo).v) | (uint32_t(__float2bfloat16(hi).v) << 16); }
inline float2 unpack_bf16x2(uint32_t u) {
  return make_float2(__bfloat162float(__nv_bfloat16{uint16_t(u & 0xffffu)}), __bfloat162float(__nv_bfloat16{uint16_t(u >> 16)}));
}
inline uint32_t pack_f16x2(float lo, float hi) { return uint32_t(f32_to_f16(lo)) | (uint32_t(f32_to_f16(hi)) << 16); }
inline float2 unpack_f16x2(uint32_t u) { return make_float2(f16_to_f32(uint16_t(u & 0xffffu)), f16_to_f32(uint16_t(u >> 16))); }

// ---- warp reductions (shuffle trees, as the device versions) and programmatic dependent launch (no-ops: launches of a
// simulated rank are already sequential)
template <typename T>
inline T warp_reduce_sum(T v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <typename T>
inline T warp_reduce_max(T v) {
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
inline void gdc_wait() {}
inline void gdc_launch_dependents() {}

// ---- NVLS emulation: `mc` is an address inside the node's multicast window
inline char* mc_copy(const void* mc, int rank) {
  return sim_mc.heap + size_t(rank) * sim_mc.bytes + (static_cast<const char*>(mc) - sim_mc.mc_base);
}
inline void multimem_st_v4(void* mc, const int4& v) {
  for (int r = 0; r < sim_mc.world; ++r) st_relaxed_sys_v4(mc_copy(mc, r), v);  // also carries LL packets
}
inline void multimem_st_b32(void* mc, uint32_t v) {
  sim_jitter();
  sim_straggle(sim_rank);
  for (int r = 0; r < sim_mc.world; ++r) __atomic_store_n(reinterpret_cast<uint32_t*>(mc_copy(mc, r)), v, kRelease);
}
inline void multimem_red_add_u32(void* mc, uint32_t v) {
  for (int r = 0; r < sim_mc.world; ++r) __atomic_fetch_add(reinterpret_cast<uint32_t*>(mc_copy(mc, r)), v, kRelease);
}
template <typename Unpack, typename PackF>
inline int4 mc_reduce16(const void* mc, Unpack unpack, PackF pack) {
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int r = 0; r < sim_mc.world; ++r) {
    const int4 v = ld_v4(mc_copy(mc, r));
    const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
    for (int i = 0; i < 4; ++i) {
      const float2 f = unpack(u[i]);
      acc[2 * i] += f.x;
      acc[2 * i + 1] += f.y;
    }
  }
  int4 out;
  uint32_t* o = reinterpret_cast<uint32_t*>(&out);
  for (int i = 0; i < 4; ++i) o[i] = pack(acc[2 * i], acc[2 * i + 1]);
  return out;
}
inline int4 multimem_ld_reduce_bf16x8(const void* mc) { return mc_reduce16(mc, unpack_bf16x2, pack_bf16x2); }
inline int4 multimem_ld_reduce_f16x8(const void* mc) { return mc_reduce16(mc, unpack_f16x2, pack_f16x2); }
inline int4 multimem_ld_reduce_f32x4(const void* mc) {
  float acc[4] = {0, 0, 0, 0};
  for (int r = 0; r < sim_mc.world; ++r) {
    const int4 v = ld_v4(mc_copy(mc, r));
    acc[0] += __int_as_float(v.x);
    acc[1] += __int_as_float(v.y);
    acc[2] += __int_as_float(v.z);
    acc[3] += __int_as_float(v.w);
  }
  return make_int4(__float_as_int(acc[0]), __float_as_int(acc[1]), __float_as_int(acc[2]), __float_as_int(acc[3]));
}

}  // namespace ptx
}  // namespace tdist
inline uint16_t sim_f32_to_f16(float f) { return tdist::ptx::f32_to_f16(f); }
inline float sim_f16_to_f32(uint16_t h) { return tdist::ptx::f16_to_f32(h); }
