// link_probe.cu — bandwidth probes of the data paths the fused kernels are built from (no counterpart in the reference;
// its comm_perf_model.py takes link speeds from sysfs / tables). Each probe is a pure data mover without cross-rank
// waits, so all ranks can be timed back to back with CUDA events; triton_dist/benchmark/bench_link_probe.py runs the
// sweep and writes the table that sizes push CTAs, reducer warps and loads in flight in gemm_sm100.cuh / collectives.cu.
//
//   mode 0  local copy                      ld.nc -> st                         (HBM reference)
//   mode 1  NVLS broadcast push             ld.nc local -> multimem.st          (AG+GEMM push, all-gather)
//   mode 2  unicast push to the W-1 peers   ld.nc local -> st peer x (W-1)
//   mode 3  pull from the W-1 peers         ld peer -> st local
//   mode 4  NVLS pull-reduce                multimem.ld_reduce -> st local      (GEMM+RS, reduce-scatter, all-reduce)
//   mode 5  bulk-copy ring to the multicast address   cp.async.bulk g2s local, s2g -> multicast (one thread per CTA)
//   mode 6  bulk-copy ring pulling from a peer        cp.async.bulk g2s peer, s2g -> local     (one thread per CTA)
#include "common.cuh"

namespace tdist {

struct ProbeArgs {
  SymmCtx c;
  const char* src;  // local source (modes 0-2, 5) — plain device memory
  char* symm;       // symmetric buffer: [W][bytes] (push targets / pull + reduce sources)
  char* out;        // local destination (modes 0, 3, 4, 6)
  size_t bytes;     // bytes per rank (multiple of 16; bulk modes: multiple of 32 KB)
  int mode;
};

template <int U>
__global__ void __launch_bounds__(512) link_probe_kernel(ProbeArgs a) {
  const SymmCtx& c = a.c;
  const int W = c.world;
  const size_t n_vec = a.bytes / 16;
  const size_t tid = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
  const size_t nthr = size_t(gridDim.x) * blockDim.x;
  if (a.mode == 0 || a.mode == 1 || a.mode == 2) {
    char* mc = mc_ptr(c, a.symm) + size_t(c.rank) * a.bytes;
    for (size_t i0 = tid; i0 < n_vec; i0 += nthr * U) {
      int4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * nthr < n_vec) v[u] = ptx::ld_nc_v4(a.src + (i0 + u * nthr) * 16);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = i0 + u * nthr;
        if (i >= n_vec) continue;
        if (a.mode == 0) {
          ptx::st_v4(a.out + i * 16, v[u]);
        } else if (a.mode == 1) {
          ptx::multimem_st_v4(mc + i * 16, v[u]);
        } else {
#pragma unroll 1
          for (int p = 1; p < W; ++p)
            ptx::st_v4(symm_at(c, a.symm, (c.rank + p) % W) + size_t(c.rank) * a.bytes + i * 16, v[u]);
        }
      }
    }
  } else if (a.mode == 3) {
    // all W-1 peer shards as one index space, peers interleaved so that every link direction is busy all the time
    const size_t total = n_vec * (W - 1);
    for (size_t i0 = tid; i0 < total; i0 += nthr * U) {
      int4 v[U];
      size_t off[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = i0 + u * nthr;
        if (i < total) {
          const int peer = (c.rank + 1 + int(i % (W - 1))) % W;
          off[u] = size_t(peer) * a.bytes + (i / (W - 1)) * 16;
          v[u] = ptx::ld_v4(symm_at(c, a.symm, peer) + off[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * nthr < total) ptx::st_v4(a.out + off[u], v[u]);
    }
  } else if (a.mode == 4) {
    const char* mc = mc_ptr(c, a.symm) + size_t(c.rank) * a.bytes;
    for (size_t i0 = tid; i0 < n_vec; i0 += nthr * U) {
      int4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * nthr < n_vec) v[u] = ptx::multimem_ld_reduce_bf16x8(mc + (i0 + u * nthr) * 16);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * nthr < n_vec) ptx::st_v4(a.out + (i0 + u * nthr) * 16, v[u]);
    }
  }
}

// one thread per CTA drives a ring of 32 KB bulk copies: global -> shared -> global
__global__ void __launch_bounds__(32) link_probe_bulk_kernel(ProbeArgs a) {
  constexpr int kChunk = 32 * 1024, kSlots = 6;
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSlots * kChunk);
  if (threadIdx.x == 0) {
    for (int i = 0; i < kSlots; ++i) ptx::mbar_init(bars + i, 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const SymmCtx& c = a.c;
  const int W = c.world;
  const size_t chunks_per_rank = a.bytes / kChunk;
  // mode 5: my shard -> multicast; mode 6: the shards of the W-1 peers -> local out
  const size_t total = a.mode == 5 ? chunks_per_rank : chunks_per_rank * (W - 1);
  auto src_of = [&](size_t k) -> const char* {
    if (a.mode == 5) return a.src + k * kChunk;
    const int peer = (c.rank + 1 + int(k % (W - 1))) % W;
    return symm_at(c, a.symm, peer) + size_t(peer) * a.bytes + (k / (W - 1)) * kChunk;
  };
  auto dst_of = [&](size_t k) -> char* {
    if (a.mode == 5) return mc_ptr(c, a.symm) + size_t(c.rank) * a.bytes + k * kChunk;
    const int peer = (c.rank + 1 + int(k % (W - 1))) % W;
    return a.out + size_t(peer) * a.bytes + (k / (W - 1)) * kChunk;
  };
  uint32_t bar_phase = 0;
  size_t issued = 0;  // index into this CTA's chunk list k = blockIdx.x + n * gridDim.x
  auto chunk_id = [&](size_t n) { return blockIdx.x + n * gridDim.x; };
  const size_t mine = total > blockIdx.x ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  for (; issued < mine && issued < size_t(kSlots); ++issued) {
    ptx::mbar_arrive_expect_tx(bars + issued, kChunk);
    ptx::bulk_g2s(smem + issued * kChunk, src_of(chunk_id(issued)), kChunk, bars + issued);
  }
  for (size_t n = 0; n < mine; ++n) {
    const int slot = int(n % kSlots);
    ptx::mbar_wait(bars + slot, (bar_phase >> slot) & 1);
    bar_phase ^= 1u << slot;
    ptx::bulk_s2g(dst_of(chunk_id(n)), smem + slot * kChunk, kChunk);
    ptx::bulk_commit();
    if (n >= 1 && issued < mine) {
      ptx::bulk_wait_read<1>();  // the previous chunk's shared memory has been read: refill its slot
      const int rs = int((n - 1) % kSlots);
      ptx::mbar_arrive_expect_tx(bars + rs, kChunk);
      ptx::bulk_g2s(smem + rs * kChunk, src_of(chunk_id(issued)), kChunk, bars + rs);
      ++issued;
    }
  }
  ptx::bulk_wait<0>();
}

}  // namespace tdist

using namespace tdist;

extern "C" int tdist_link_probe(const tdist_symm_ctx* ctx, int mode, const void* src, void* symm, void* out, size_t bytes,
                                int grid, int threads, int unroll, void* stream) {
  ProbeArgs a{to_ctx(ctx), static_cast<const char*>(src), static_cast<char*>(symm), static_cast<char*>(out), bytes, mode};
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (bytes % 16 != 0 || grid < 1) {
    set_error("link_probe: bytes must be a multiple of 16 and grid positive");
    return -1;
  }
  if ((mode == 1 || mode == 4 || mode == 5) && a.c.mc_base == nullptr) {
    set_error("link_probe: this mode needs a multicast mapping");
    return -1;
  }
  if (mode == 5 || mode == 6) {
    if (bytes % (32 * 1024) != 0) {
      set_error("link_probe: bulk modes need a multiple of 32 KB");
      return -1;
    }
    constexpr int kSmem = 6 * 32 * 1024 + 64;
    static bool attr_set = false;
    if (!attr_set) {
      if (cudaFuncSetAttribute(link_probe_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem) != cudaSuccess) {
        set_error("link_probe: cannot raise the shared memory limit");
        return -1;
      }
      attr_set = true;
    }
    link_probe_bulk_kernel<<<grid, 32, kSmem, s>>>(a);
    return check_launch("link_probe_bulk");
  }
  if (mode < 0 || mode > 4 || threads < 32 || threads > 512) {
    set_error("link_probe: unknown mode or bad thread count");
    return -1;
  }
  switch (unroll) {
    case 4: link_probe_kernel<4><<<grid, threads, 0, s>>>(a); break;
    case 8: link_probe_kernel<8><<<grid, threads, 0, s>>>(a); break;
    case 16: link_probe_kernel<16><<<grid, threads, 0, s>>>(a); break;
    default: set_error("link_probe: unroll must be 4, 8 or 16"); return -1;
  }
  return check_launch("link_probe");
}
