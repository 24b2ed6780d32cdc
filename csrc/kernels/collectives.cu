// collectives.cu — barrier, AllReduce, AllGather, ReduceScatter and P2P copy
// kernels over the symmetric heap (NVLink P2P stores/loads and NVLS multimem).
//
// Reference counterparts (behaviour, not code): kernels/nvidia/allreduce.py
// K16-K22 (:214-682), low_latency_allgather.py K28-K36 (:47-777),
// reduce_scatter.py K12-K15 (:182-742), common_ops.py K2/K3 (:141-209),
// kernels/nvidia/p2p.py K58 (:30-85).
//
// B200 design notes
//  * NVSwitch is flat, so every ring / NUMA / 2-D / 3-D variant of the
//    reference collapses to a one-hop exchange.
//  * Cross-GPU synchronisation is per CTA: CTA b of every rank owns the same
//    slice of the message, so "all ranks finished slice b" needs only a W-wide
//    flag row per CTA and no grid-wide barrier.
//  * Flags carry a monotone phase (never reset) and staging buffers are
//    double-buffered by phase parity, which removes the trailing barrier of
//    the reference's kernels and keeps every launch CUDA-graph capturable.
#include "common.cuh"

namespace tdist {

// per-CTA cross-rank barrier on this CTA's flag row
__device__ __forceinline__ void cta_barrier_all(const SymmCtx& c, uint32_t* flags, uint32_t phase) {
  barrier_all_block(c, flags + blockIdx.x * TDIST_MAX_RANKS, phase);
}

// Bulk loops keep kIlp independent 16-byte loads per thread in flight before the first store: a single load per
// thread and iteration left the large-message kernels latency-bound (64 MB reduce-scatter 214 us vs NCCL 139 us; the
// link probe reaches the NVLS rate with 8 loads in flight on 148 CTAs, profiles/README.md).
constexpr int kIlp = 8;
template <typename LoadF, typename StoreF>
__device__ __forceinline__ void bulk_loop(size_t beg, size_t end, LoadF load, StoreF store) {
  const size_t stride = blockDim.x;
  size_t i = beg + threadIdx.x;
  for (; i + (kIlp - 1) * stride < end; i += kIlp * stride) {
    int4 v[kIlp];
#pragma unroll
    for (int u = 0; u < kIlp; ++u) v[u] = load(i + u * stride);
#pragma unroll
    for (int u = 0; u < kIlp; ++u) store(i + u * stride, v[u]);
  }
  for (; i < end; i += stride) store(i, load(i));
}

// ------------------------------------------------------------------ barrier
__global__ void barrier_all_kernel(SymmCtx c, uint32_t* flags, PhaseState ps) {
  const uint32_t base = phase_begin(ps);
  barrier_all_block(c, flags, base + 1);
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ AllReduce: one-shot push
// ws layout: [2 parity][W slots][slot_vecs] int4. Each rank stores its slice
// into slot `rank` of every peer, then sums the W slots locally.
template <typename Tag>
__global__ void __launch_bounds__(512) ar_one_shot_push_kernel(SymmCtx c, const int4* __restrict__ in, int4* out,
                                                               int4* ws, size_t slot_vecs, uint32_t* flags,
                                                               PhaseState ps, size_t n_vec) {
  const uint32_t base = phase_begin(ps);
  const uint32_t phase = base + 1;
  int4* my_ws = ws + ((base >> 1) & 1) * c.world * slot_vecs;
  const size_t per_cta = (n_vec + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, n_vec);
  for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    const int4 v = ptx::ld_nc_v4(in + i);
#pragma unroll 1
    for (int p = 0; p < c.world; ++p) {
      const int peer = (c.rank + p) % c.world;
      ptx::st_v4(symm_at(c, my_ws, peer) + c.rank * slot_vecs + i, v);
    }
  }
  cta_barrier_all(c, flags, phase);
  for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int s = 0; s < c.world; ++s) Pack<Tag>::add(acc, ptx::ld_v4(my_ws + s * slot_vecs + i));
    out[i] = Pack<Tag>::cvt(acc);
  }
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ AllReduce: two-shot push
// ws layout per parity: [W slots][shard_vecs] scatter area followed by
// [n_vec] result area. Rank r owns the r-th 1/W of every CTA slice.
template <typename Tag>
__global__ void __launch_bounds__(512) ar_two_shot_push_kernel(SymmCtx c, const int4* __restrict__ in, int4* out,
                                                               int4* ws, size_t parity_vecs, size_t shard_cap,
                                                               uint32_t* flags, PhaseState ps, size_t n_vec) {
  const uint32_t base = phase_begin(ps);
  const int W = c.world;
  int4* scat = ws + ((base >> 1) & 1) * parity_vecs;  // [W][shard_cap]; shard_cap bounds one rank's share
  int4* res = scat + W * shard_cap;
  const size_t per_cta = (n_vec + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, n_vec);
  const size_t len = end > beg ? end - beg : 0;
  const size_t per_rank = (len + W - 1) / W;
  // position of (cta slice, owner-local index j) inside an owner's slot
  const size_t slot_off = blockIdx.x * ((per_cta + W - 1) / W);
  // phase 1: scatter my contribution of owner o's sub-slice into o's slot [rank]
#pragma unroll 1
  for (int p = 0; p < W; ++p) {
    const int o = (c.rank + p) % W;
    const size_t sb = min(beg + o * per_rank, end), se = min(sb + per_rank, end);
    int4* dst = symm_at(c, scat, o) + c.rank * shard_cap + slot_off;
    for (size_t i = sb + threadIdx.x; i < se; i += blockDim.x) ptx::st_v4(dst + (i - sb), ptx::ld_nc_v4(in + i));
  }
  cta_barrier_all(c, flags, base + 1);
  // phase 2: reduce my sub-slice and broadcast the result to every rank's result area
  {
    const size_t sb = min(beg + c.rank * per_rank, end), se = min(sb + per_rank, end);
    for (size_t i = sb + threadIdx.x; i < se; i += blockDim.x) {
      float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int s = 0; s < W; ++s) Pack<Tag>::add(acc, ptx::ld_v4(scat + s * shard_cap + slot_off + (i - sb)));
      const int4 r = Pack<Tag>::cvt(acc);
#pragma unroll 1
      for (int p = 0; p < W; ++p) ptx::st_v4(symm_at(c, res, (c.rank + p) % W) + i, r);
    }
  }
  cta_barrier_all(c, flags, base + 2);
  for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) out[i] = ptx::ld_v4(res + i);
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ AllReduce: NVLS one-shot
// symm: [2 parity][n_vec]. copy-in, per-CTA barrier, multimem.ld_reduce -> out.
template <typename Tag>
__global__ void __launch_bounds__(512) ar_one_shot_mc_kernel(SymmCtx c, const int4* __restrict__ in, int4* out,
                                                             int4* symm, size_t parity_vecs, uint32_t* flags,
                                                             PhaseState ps, size_t n_vec, int in_place) {
  const uint32_t base = phase_begin(ps);
  const uint32_t phase = base + 1;
  int4* buf = in_place ? symm : symm + ((base >> 1) & 1) * parity_vecs;
  const size_t per_cta = (n_vec + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, n_vec);
  if (!in_place)
    bulk_loop(beg, end, [&](size_t i) { return ptx::ld_nc_v4(in + i); }, [&](size_t i, const int4& v) { ptx::st_v4(buf + i, v); });
  cta_barrier_all(c, flags, phase);
  const int4* mc = mc_ptr(c, buf);
  bulk_loop(beg, end, [&](size_t i) { return Pack<Tag>::mc_ld_reduce(mc + i); }, [&](size_t i, const int4& v) { out[i] = v; });
  if (in_place) cta_barrier_all(c, flags, phase + 1);  // nobody may overwrite symm before all ranks read it
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ AllReduce: NVLS two-shot
// each rank ld_reduces its 1/W of every CTA slice and multimem.st's the sum
// back in place on all ranks.
template <typename Tag>
__global__ void __launch_bounds__(512) ar_two_shot_mc_kernel(SymmCtx c, const int4* __restrict__ in, int4* out,
                                                             int4* symm, size_t parity_vecs, uint32_t* flags,
                                                             PhaseState ps, size_t n_vec, int in_place) {
  const uint32_t base = phase_begin(ps);
  const int W = c.world;
  int4* buf = in_place ? symm : symm + ((base >> 1) & 1) * parity_vecs;
  const size_t per_cta = (n_vec + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, n_vec);
  const size_t len = end > beg ? end - beg : 0;
  const size_t per_rank = (len + W - 1) / W;
  if (!in_place)
    bulk_loop(beg, end, [&](size_t i) { return ptx::ld_nc_v4(in + i); }, [&](size_t i, const int4& v) { ptx::st_v4(buf + i, v); });
  cta_barrier_all(c, flags, base + 1);
  {
    int4* mc = mc_ptr(c, buf);
    const size_t sb = min(beg + c.rank * per_rank, end), se = min(sb + per_rank, end);
    bulk_loop(sb, se, [&](size_t i) { return Pack<Tag>::mc_ld_reduce(mc + i); },
              [&](size_t i, const int4& v) { ptx::multimem_st_v4(mc + i, v); });
  }
  cta_barrier_all(c, flags, base + 2);
  if (out != buf)
    bulk_loop(beg, end, [&](size_t i) { return ptx::ld_v4(buf + i); }, [&](size_t i, const int4& v) { out[i] = v; });
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ AllGather
// symm: [W][shard_vecs] on every rank (slot r = rank r's shard).
// mode 0 pull: publish own shard locally, barrier, read peers' slots.
// mode 1 push: store own shard into slot `rank` of every peer, barrier.
// mode 2 multimem push: one multimem.st per 16 B reaches all ranks.
__global__ void __launch_bounds__(512) allgather_kernel(SymmCtx c, const int4* __restrict__ in, int4* symm,
                                                        int4* out, size_t shard_vecs, uint32_t* flags,
                                                        PhaseState ps, int mode) {
  const uint32_t base = phase_begin(ps);
  const int W = c.world;
  const size_t per_cta = (shard_vecs + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, shard_vecs);
  int4* mine = symm + c.rank * shard_vecs;
  if (mode == 0) {
    if (in != mine)
      for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) ptx::st_v4(mine + i, ptx::ld_nc_v4(in + i));
    cta_barrier_all(c, flags, base + 1);
#pragma unroll 1
    for (int p = 1; p < W; ++p) {
      const int peer = (c.rank + p) % W;
      const int4* src = symm_at(c, symm, peer) + peer * shard_vecs;
      for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) ptx::st_v4(symm + peer * shard_vecs + i, ptx::ld_v4(src + i));
    }
    cta_barrier_all(c, flags, base + 2);  // peers may not overwrite their slot while I still read it
  } else if (mode == 1) {
    for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
      const int4 v = ptx::ld_nc_v4(in + i);
#pragma unroll 1
      for (int p = 0; p < W; ++p) ptx::st_v4(symm_at(c, symm, (c.rank + p) % W) + c.rank * shard_vecs + i, v);
    }
    cta_barrier_all(c, flags, base + 1);
  } else {
    int4* mc = mc_ptr(c, symm) + c.rank * shard_vecs;
    for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) ptx::multimem_st_v4(mc + i, ptx::ld_nc_v4(in + i));
    cta_barrier_all(c, flags, base + 1);
  }
  if (out != nullptr && out != symm) {
    for (int s = 0; s < W; ++s)
      for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x)
        out[s * shard_vecs + i] = ptx::ld_v4(symm + s * shard_vecs + i);
  }
  phase_end(ps, base, 2);
}

// LL all-gather: 16-byte packets {d0, flag, d1, flag}; the receiver spins on
// the packet itself, so a message costs one NVLink one-way trip and no fence.
// llbuf: [2 parity][W][shard_u64] packets (16 B each). in/out are 8-byte units.
__global__ void __launch_bounds__(512) allgather_ll_kernel(SymmCtx c, const uint2* __restrict__ in, int4* llbuf,
                                                           uint2* out, size_t shard_u64, size_t parity_pkts,
                                                           PhaseState ps, int use_mc) {
  const uint32_t base = phase_begin(ps);
  const uint32_t flag = base + 1;
  const int W = c.world;
  int4* buf = llbuf + ((base >> 1) & 1) * parity_pkts;  // fixed distance: half the buffer, not a size of this call
  const size_t per_cta = (shard_u64 + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, shard_u64);
  for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    const uint2 d = in[i];
    int4 pkt;
    pkt.x = d.x;
    pkt.y = flag;
    pkt.z = d.y;
    pkt.w = flag;
    if (use_mc) {
      ptx::multimem_st_v4(mc_ptr(c, buf) + c.rank * shard_u64 + i, pkt);
    } else {
#pragma unroll 1
      for (int p = 0; p < W; ++p) ptx::st_relaxed_sys_v4(symm_at(c, buf, (c.rank + p) % W) + c.rank * shard_u64 + i, pkt);
    }
  }
  for (int s = 0; s < W; ++s) {
    for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
      int4 pkt;
      do {
        pkt = ptx::ld_volatile_v4(buf + s * shard_u64 + i);
      } while (static_cast<uint32_t>(pkt.y) != flag || static_cast<uint32_t>(pkt.w) != flag);
      out[s * shard_u64 + i] = make_uint2(pkt.x, pkt.z);
    }
  }
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ ReduceScatter
// in: [W][seg_vecs] (segment s goes to rank s); scat (symm): [2 parity][W][seg_vecs];
// out: [seg_vecs] = sum over ranks of their segment `rank`.
template <typename Tag>
__global__ void __launch_bounds__(512) reduce_scatter_kernel(SymmCtx c, const int4* __restrict__ in, int4* scat,
                                                             int4* out, size_t seg_vecs, size_t parity_vecs,
                                                             uint32_t* flags, PhaseState ps) {
  const uint32_t base = phase_begin(ps);
  const uint32_t phase = base + 1;
  const int W = c.world;
  int4* buf = scat + ((base >> 1) & 1) * parity_vecs;  // fixed distance between the parity halves
  const size_t per_cta = (seg_vecs + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, seg_vecs);
#pragma unroll 1
  for (int p = 1; p < W; ++p) {
    const int o = (c.rank + p) % W;
    int4* dst = symm_at(c, buf, o) + c.rank * seg_vecs;
    for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) ptx::st_v4(dst + i, ptx::ld_nc_v4(in + o * seg_vecs + i));
  }
  cta_barrier_all(c, flags, phase);
  for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    Pack<Tag>::add(acc, ptx::ld_nc_v4(in + c.rank * seg_vecs + i));
    for (int p = 1; p < W; ++p) Pack<Tag>::add(acc, ptx::ld_v4(buf + ((c.rank + p) % W) * seg_vecs + i));
    out[i] = Pack<Tag>::cvt(acc);
  }
  phase_end(ps, base, 2);
}

// NVLS reduce-scatter: every rank copies its W segments into the symmetric buffer (same layout everywhere), CTA b of
// all ranks meet, then each rank sums ITS segment across ranks in the switch. 1/W of the data enters each GPU instead
// of (W-1)/W, and nothing is reduced on the SMs.
template <typename Tag>
__global__ void __launch_bounds__(512) reduce_scatter_mc_kernel(SymmCtx c, const int4* __restrict__ in, int4* scat,
                                                                int4* out, size_t seg_vecs, size_t parity_vecs,
                                                                uint32_t* flags, PhaseState ps) {
  const uint32_t base = phase_begin(ps);
  const int W = c.world;
  int4* buf = scat + ((base >> 1) & 1) * parity_vecs;
  const size_t per_cta = (seg_vecs + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, seg_vecs);
#pragma unroll 1
  for (int s = 0; s < W; ++s) {
    const int4* src = in + s * seg_vecs;
    int4* dst = buf + s * seg_vecs;
    bulk_loop(beg, end, [&](size_t i) { return ptx::ld_nc_v4(src + i); }, [&](size_t i, const int4& v) { ptx::st_v4(dst + i, v); });
  }
  cta_barrier_all(c, flags, base + 1);
  const int4* mc = mc_ptr(c, buf) + c.rank * seg_vecs;
  bulk_loop(beg, end, [&](size_t i) { return Pack<Tag>::mc_ld_reduce(mc + i); }, [&](size_t i, const int4& v) { out[i] = v; });
  phase_end(ps, base, 2);
}

// out[i] = sum_s in[s][i]  (the reference's ring_reduce, reduce_scatter.py:673-742; order: rank+1 ... rank)
template <typename Tag>
__global__ void __launch_bounds__(512) slot_reduce_kernel(const int4* __restrict__ in, int4* out, size_t seg_vecs,
                                                          int slots, int begin) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < seg_vecs; i += size_t(gridDim.x) * blockDim.x) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int s = 0; s < slots; ++s) Pack<Tag>::add(acc, ptx::ld_nc_v4(in + ((begin + 1 + s) % slots) * seg_vecs + i));
    out[i] = Pack<Tag>::cvt(acc);
  }
}

// ------------------------------------------------------------------ P2P copy (pipeline-parallel activations)
__global__ void __launch_bounds__(512) p2p_copy_kernel(int4* dst, const int4* src, size_t n_vec) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n_vec; i += size_t(gridDim.x) * blockDim.x)
    ptx::st_v4(dst + i, ptx::ld_v4(src + i));
}

// one-thread signal helpers usable from a stream (graph-capturable alternative to stream mem-ops)
__global__ void signal_set_kernel(uint32_t* flag, uint32_t value, int add) {
  ptx::fence_acq_rel_sys();
  if (add)
    ptx::red_add_release_sys(flag, value);
  else
    ptx::st_release_sys(flag, value);
}
__global__ void signal_wait_kernel(const uint32_t* flag, uint32_t value, int cmp) {
  signal_wait_until(flag, static_cast<CmpOp>(cmp), value);
}

static inline int pick_grid(size_t n_vec, int max_ctas, int threads) {
  size_t want = (n_vec + size_t(threads) * 4 - 1) / (size_t(threads) * 4);
  if (want < 1) want = 1;
  if (want > size_t(max_ctas)) want = max_ctas;
  return static_cast<int>(want);
}

}  // namespace tdist

using namespace tdist;

extern "C" {

int tdist_barrier_all(const tdist_symm_ctx* ctx, uint32_t* flags, uint32_t* phase, uint32_t* done, void* stream) {
  barrier_all_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(to_ctx(ctx), flags, PhaseState{phase, done});
  return check_launch("barrier_all");
}

int tdist_allreduce(const tdist_symm_ctx* ctx, int method, int dtype, const void* in, void* out, size_t nbytes,
                    void* ws, size_t ws_bytes, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_ctas,
                    int in_place, void* stream) {
  if (nbytes % 16 != 0) {
    set_error("allreduce: nbytes must be a multiple of 16");
    return -1;
  }
  const size_t n_vec = nbytes / 16;
  const int threads = 512;
  if (max_ctas > TDIST_MAX_CTAS) max_ctas = TDIST_MAX_CTAS;
  int grid = pick_grid(n_vec, max_ctas, threads);
  const SymmCtx c = to_ctx(ctx);
  const PhaseState ps{phase, done};
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int W = c.world;
  if ((method == TDIST_AR_ONE_SHOT_MC || method == TDIST_AR_TWO_SHOT_MC) && c.mc_base == nullptr) {
    set_error("allreduce: multimem method requested but the heap has no multicast mapping");
    return -1;
  }
  TDIST_DISPATCH_DTYPE(dtype, {
    switch (method) {
      case TDIST_AR_ONE_SHOT: {
        const size_t slot_vecs = ws_bytes / 16 / 2 / W;
        if (slot_vecs < n_vec) {
          set_error("allreduce one_shot: workspace too small (need 2*W*nbytes)");
          return -1;
        }
        ar_one_shot_push_kernel<Tag><<<grid, threads, 0, s>>>(c, static_cast<const int4*>(in), static_cast<int4*>(out),
                                                               static_cast<int4*>(ws), slot_vecs, flags, ps, n_vec);
        break;
      }
      case TDIST_AR_TWO_SHOT: {
        const size_t shard_cap = (n_vec + W - 1) / W + 2 * size_t(grid);
        // the two parity halves are half the workspace apart whatever this call's size is: calls of different sizes
        // must not overlap the half a slower rank is still reading
        const size_t parity_vecs = ws_bytes / 16 / 2;
        if (parity_vecs < W * shard_cap + n_vec) {
          set_error("allreduce two_shot: workspace too small");
          return -1;
        }
        ar_two_shot_push_kernel<Tag><<<grid, threads, 0, s>>>(c, static_cast<const int4*>(in), static_cast<int4*>(out),
                                                               static_cast<int4*>(ws), parity_vecs, shard_cap, flags, ps,
                                                               n_vec);
        break;
      }
      case TDIST_AR_ONE_SHOT_MC: {
        const size_t parity_vecs = in_place ? 0 : ws_bytes / 16 / 2;
        if (!in_place && parity_vecs < n_vec) {
          set_error("allreduce one_shot_multimem: workspace too small (need 2*nbytes)");
          return -1;
        }
        ar_one_shot_mc_kernel<Tag><<<grid, threads, 0, s>>>(c, static_cast<const int4*>(in), static_cast<int4*>(out),
                                                             static_cast<int4*>(ws), parity_vecs, flags, ps, n_vec,
                                                             in_place);
        break;
      }
      case TDIST_AR_TWO_SHOT_MC: {
        const size_t parity_vecs = in_place ? 0 : ws_bytes / 16 / 2;
        if (!in_place && parity_vecs < n_vec) {
          set_error("allreduce two_shot_multimem: workspace too small (need 2*nbytes)");
          return -1;
        }
        ar_two_shot_mc_kernel<Tag><<<grid, threads, 0, s>>>(c, static_cast<const int4*>(in), static_cast<int4*>(out),
                                                             static_cast<int4*>(ws), parity_vecs, flags, ps, n_vec,
                                                             in_place);
        break;
      }
      default:
        set_error("allreduce: unknown method");
        return -1;
    }
  });
  return check_launch("allreduce");
}

int tdist_allgather(const tdist_symm_ctx* ctx, int mode, const void* in, void* symm, void* out, size_t shard_bytes,
                    size_t symm_bytes, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_ctas, void* stream) {
  const SymmCtx c = to_ctx(ctx);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const PhaseState ps{phase, done};
  if (max_ctas > TDIST_MAX_CTAS) max_ctas = TDIST_MAX_CTAS;
  if (mode == TDIST_AG_LL || mode == TDIST_AG_LL_MC) {
    if (shard_bytes % 8 != 0) {
      set_error("allgather LL: shard bytes must be a multiple of 8");
      return -1;
    }
    if (mode == TDIST_AG_LL_MC && c.mc_base == nullptr) {
      set_error("allgather LL multimem: no multicast mapping");
      return -1;
    }
    const size_t n = shard_bytes / 8;
    const size_t parity_pkts = symm_bytes / 16 / 2;
    if (parity_pkts < size_t(c.world) * n) {
      set_error("allgather LL: packet buffer too small (need 2 * W * shard_bytes * 2 bytes)");
      return -1;
    }
    int grid = pick_grid(n, max_ctas, 512);
    allgather_ll_kernel<<<grid, 512, 0, s>>>(c, static_cast<const uint2*>(in), static_cast<int4*>(symm),
                                             static_cast<uint2*>(out), n, parity_pkts, ps, mode == TDIST_AG_LL_MC);
    return check_launch("allgather_ll");
  }
  if (shard_bytes % 16 != 0) {
    set_error("allgather: shard bytes must be a multiple of 16");
    return -1;
  }
  if (mode == TDIST_AG_PUSH_MC && c.mc_base == nullptr) {
    set_error("allgather multimem: no multicast mapping");
    return -1;
  }
  const size_t n = shard_bytes / 16;
  int grid = pick_grid(n, max_ctas, 512);
  allgather_kernel<<<grid, 512, 0, s>>>(c, static_cast<const int4*>(in), static_cast<int4*>(symm),
                                        static_cast<int4*>(out), n, flags, ps, mode);
  return check_launch("allgather");
}

int tdist_reduce_scatter(const tdist_symm_ctx* ctx, int dtype, const void* in, void* scatter_ws, size_t ws_bytes,
                         void* out, size_t seg_bytes, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_ctas,
                         void* stream) {
  if (seg_bytes % 16 != 0) {
    set_error("reduce_scatter: segment bytes must be a multiple of 16");
    return -1;
  }
  const SymmCtx c = to_ctx(ctx);
  const size_t n = seg_bytes / 16;
  const size_t parity_vecs = ws_bytes / 16 / 2;
  if (parity_vecs < size_t(c.world) * n) {
    set_error("reduce_scatter: workspace too small (need 2 * W * seg_bytes)");
    return -1;
  }
  if (max_ctas > TDIST_MAX_CTAS) max_ctas = TDIST_MAX_CTAS;
  // NVLS from 1 MB per segment (below, the barrier + copy-in costs more than the one-hop scatter saves)
  const bool nvls = c.mc_base != nullptr && c.world > 1 && seg_bytes >= (512u << 10) && getenv("TDIST_RS_NO_NVLS") == nullptr;
  if (nvls && max_ctas < 64) max_ctas = TDIST_MAX_CTAS < 148 ? TDIST_MAX_CTAS : 148;  // the pull needs the whole GPU's loads in flight
  int grid = pick_grid(n, max_ctas, 512);
  TDIST_DISPATCH_DTYPE(dtype, {
    if (nvls)
      reduce_scatter_mc_kernel<Tag><<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(
          c, static_cast<const int4*>(in), static_cast<int4*>(scatter_ws), static_cast<int4*>(out), n, parity_vecs, flags,
          PhaseState{phase, done});
    else
      reduce_scatter_kernel<Tag><<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(
          c, static_cast<const int4*>(in), static_cast<int4*>(scatter_ws), static_cast<int4*>(out), n, parity_vecs, flags,
          PhaseState{phase, done});
  });
  return check_launch("reduce_scatter");
}

int tdist_slot_reduce(int dtype, const void* in, void* out, size_t seg_bytes, int slots, int begin, int max_ctas,
                      void* stream) {
  if (seg_bytes % 16 != 0) {
    set_error("slot_reduce: segment bytes must be a multiple of 16");
    return -1;
  }
  const size_t n = seg_bytes / 16;
  int grid = pick_grid(n, max_ctas, 512);
  TDIST_DISPATCH_DTYPE(dtype, {
    slot_reduce_kernel<Tag><<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const int4*>(in), static_cast<int4*>(out), n, slots, begin);
  });
  return check_launch("slot_reduce");
}

int tdist_p2p_copy(void* dst, const void* src, size_t nbytes, int num_ctas, void* stream) {
  if (nbytes % 16 != 0) {
    set_error("p2p_copy: nbytes must be a multiple of 16");
    return -1;
  }
  const size_t n = nbytes / 16;
  int grid = pick_grid(n, num_ctas, 512);
  p2p_copy_kernel<<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<int4*>(dst),
                                                                      static_cast<const int4*>(src), n);
  return check_launch("p2p_copy");
}

int tdist_signal_set(uint32_t* flag, uint32_t value, int add, void* stream) {
  signal_set_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(flag, value, add);
  return check_launch("signal_set");
}
int tdist_signal_wait(const uint32_t* flag, uint32_t value, int cmp, void* stream) {
  signal_wait_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(flag, value, cmp);
  return check_launch("signal_wait");
}

}  // extern "C"
