// ep_a2a.cu — expert-parallel all-to-all over NVLink stores.
//
//  * a2a_kernel: the fixed-slot low-latency all-to-all of the reference
//    (kernels/nvidia/low_latency_all_to_all.py:35-118): tokens arrive sorted by expert, each peer gets
//    one contiguous row range, its per-expert splits and (optionally) per-token scales, double-buffered
//    by call parity; several CTAs per peer instead of the reference's single one.
//  * ep_counts / ep_dispatch / ep_combine: token-level dispatch/combine of the DeepEP-style layer
//    (kernels/nvidia/ep_a2a.py:36-334, layers/nvidia/ep_a2a_layer.py:118-248) for ONE NVSwitch node:
//    per-expert counts are all-gathered by stores, every rank derives the exact receive layout
//    (rows ordered by local expert, then source rank), tokens are written straight into the destination
//    rank's buffer at their final row, optionally quantised to fp8 e4m3 with one fp32 scale per 128
//    values on the way (the only fp8 path of the reference is tutorials/04-deepseek-infer-all2all.py:334-427);
//    combine pulls the top-k expert outputs of every token from the peers and sums them in fp32.
#include <cuda_fp8.h>

#include "common.cuh"

namespace tdist {

// ------------------------------------------------------------------ low-latency all-to-all
// grid = W * ctas_per_peer. recv_data: [2][W][max_m][row_bytes]; recv_splits: [2][W*epr]; recv_scale: [2][W][max_m] f32.
// flags: [2][W] arrival counters (monotone: every call adds ctas_per_peer to flags[parity][src] on the receiver).
__global__ void __launch_bounds__(512) a2a_kernel(SymmCtx c, const char* __restrict__ send, const int* __restrict__ splits_cumsum,
                                                  const float* __restrict__ send_scale, char* recv_data, int* recv_splits,
                                                  float* recv_scale, uint32_t* flags, PhaseState ps, int max_m, int row_bytes,
                                                  int epr, int ctas_per_peer) {
  const uint32_t base = phase_begin(ps);
  const uint32_t call = base / 2 + 1;  // 1, 2, 3, ...
  const int parity = call & 1;
  const uint32_t target = ((call + 1) / 2) * ctas_per_peer;  // calls with this parity so far, times CTAs
  const int W = c.world;
  const int peer = blockIdx.x / ctas_per_peer, part = blockIdx.x % ctas_per_peer;
  const int m_st = splits_cumsum[peer * epr], m_ed = splits_cumsum[(peer + 1) * epr];
  const size_t rows = size_t(m_ed - m_st);
  // payload slice of this CTA (16-byte units)
  const size_t vecs = rows * row_bytes / 16;
  const size_t per = (vecs + ctas_per_peer - 1) / ctas_per_peer;
  const size_t vb = min(size_t(part) * per, vecs), ve = min(vb + per, vecs);
  char* dst = symm_at(c, recv_data, peer) + ((size_t(parity) * W + c.rank) * max_m) * row_bytes;
  const char* src = send + size_t(m_st) * row_bytes;
  copy_block(dst + vb * 16, src + vb * 16, (ve - vb) * 16);
  if (part == 0) {
    int* sdst = symm_at(c, recv_splits, peer) + parity * W * epr + c.rank * epr;
    for (int e = threadIdx.x; e < epr; e += blockDim.x)
      sdst[e] = splits_cumsum[peer * epr + e + 1] - splits_cumsum[peer * epr + e];
    if (send_scale != nullptr) {
      float* fdst = symm_at(c, recv_scale, peer) + (size_t(parity) * W + c.rank) * max_m;
      for (size_t i = threadIdx.x; i < rows; i += blockDim.x) fdst[i] = send_scale[m_st + i];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    ptx::fence_acq_rel_sys();
    ptx::red_add_release_sys(symm_at(c, flags, peer) + parity * W + c.rank, 1u);
    // wait for the same slice from `peer` (every CTA waits for one source: W*ctas waits cover all sources)
    wait_ge_sys(flags + parity * W + peer, target);
  }
  __syncthreads();
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ EP dispatch, step 1: counts all-gather
// local_counts[E] = bincount(topk_ids); stored into full_counts[rank][E] on every rank. 1 CTA.
__global__ void __launch_bounds__(1024) ep_counts_kernel(SymmCtx c, const int* __restrict__ topk_ids, int num_pairs,
                                                         int num_experts, int* full_counts, uint32_t* flags, PhaseState ps) {
  extern __shared__ int sh_cnt[];
  const uint32_t base = phase_begin(ps);
  for (int e = threadIdx.x; e < num_experts; e += blockDim.x) sh_cnt[e] = 0;
  __syncthreads();
  for (int p = threadIdx.x; p < num_pairs; p += blockDim.x) {
    const int e = topk_ids[p];
    if (e >= 0 && e < num_experts) atomicAdd(&sh_cnt[e], 1);
  }
  __syncthreads();
  for (int peer = 0; peer < c.world; ++peer) {
    int* dst = symm_at(c, full_counts, peer) + c.rank * num_experts;
    for (int e = threadIdx.x; e < num_experts; e += blockDim.x) dst[e] = sh_cnt[e];
  }
  barrier_all_block(c, flags, base + 1);
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ EP dispatch, step 2: layout + token stores
// full_counts: [W src][E]. Receive layout on rank d: rows ordered by (local expert e, src s).
// Outputs (local): scatter_idx[pair] = row on the destination rank; cursors[E] zeroed scratch;
// recv_meta on every rank: recv_expert_counts[epr] and num_recv.
// fp8 mode: the row is quantised per 128 values to e4m3, payload = hidden bytes, scales to recv_scale[row][hidden/128].
template <typename T>
__global__ void __launch_bounds__(1024) ep_dispatch_kernel(SymmCtx c, const T* __restrict__ x, const int* __restrict__ topk_ids,
                                                           const int* __restrict__ full_counts, char* recv_buf,
                                                           float* recv_scale, int* recv_src_pair, int* scatter_idx, int* cursors,
                                                           int* recv_expert_counts, int* num_recv, int tokens, int topk, int hidden,
                                                           int num_experts, int fp8, uint32_t* flags, uint32_t* grid_counter,
                                                           PhaseState ps) {
  extern __shared__ int sh_off[];  // base row of (expert e, src = me) on its destination rank, for all E experts
  const uint32_t base = phase_begin(ps);
  const int W = c.world, epr = num_experts / W;
  // exact layout: off[e] = sum_{e' local to d, e' < e} sum_s cnt[s][e'] + sum_{s < me} cnt[s][e]
  for (int e = threadIdx.x; e < num_experts; e += blockDim.x) {
    const int d = e / epr;
    int off = 0;
    for (int e2 = d * epr; e2 < e; ++e2)
      for (int s = 0; s < W; ++s) off += full_counts[s * num_experts + e2];
    for (int s = 0; s < c.rank; ++s) off += full_counts[s * num_experts + e];
    sh_off[e] = off;
  }
  if (blockIdx.x == 0) {
    for (int e = threadIdx.x; e < epr; e += blockDim.x) {
      int n = 0;
      for (int s = 0; s < W; ++s) n += full_counts[s * num_experts + c.rank * epr + e];
      recv_expert_counts[e] = n;
    }
    if (threadIdx.x == 0) {
      int n = 0;
      for (int e = 0; e < epr; ++e)
        for (int s = 0; s < W; ++s) n += full_counts[s * num_experts + c.rank * epr + e];
      *num_recv = n;
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warps_per_cta = blockDim.x >> 5;
  const int num_pairs = tokens * topk;
  const int out_row_bytes = fp8 ? hidden : hidden * int(sizeof(T));
  for (int p = blockIdx.x * warps_per_cta + warp; p < num_pairs; p += gridDim.x * warps_per_cta) {
    const int e = topk_ids[p];
    if (e < 0 || e >= num_experts) {
      if (lane == 0) scatter_idx[p] = -1;
      continue;
    }
    const int d = e / epr;
    int slot = 0;
    if (lane == 0) slot = sh_off[e] + atomicAdd(&cursors[e], 1);
    slot = __shfl_sync(0xffffffffu, slot, 0);
    const T* src = x + size_t(p / topk) * hidden;
    char* dst = symm_at(c, recv_buf, d) + size_t(slot) * out_row_bytes;
    if (!fp8) {
      copy_warp(dst, src, size_t(hidden) * sizeof(T));
    } else {
      // 128 values per group: 4 lanes x 32 values? keep it simple: each lane owns 8 values, 16 lanes = one group
      float* sdst = symm_at(c, recv_scale, d) + size_t(slot) * (hidden / 128);
      for (int g0 = 0; g0 < hidden; g0 += 256) {  // a warp covers two 128-groups per iteration
        const int col = g0 + lane * 8;
        float v[8];
        float amax = 0.f;
        if (col < hidden) {
          const int4 raw = ptx::ld_nc_v4(src + col);
          const uint32_t* u = reinterpret_cast<const uint32_t*>(&raw);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float2 f;
            if (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value)
              f = ptx::unpack_bf16x2(u[i]);
            else
              f = ptx::unpack_f16x2(u[i]);
            v[2 * i] = f.x;
            v[2 * i + 1] = f.y;
            amax = fmaxf(amax, fmaxf(fabsf(f.x), fabsf(f.y)));
          }
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));  // 16-lane groups
        const float scale = amax > 0.f ? amax / 448.f : 1.f;
        if (col < hidden) {
          const float inv = 1.f / scale;
          uint2 packed;
          unsigned char* b = reinterpret_cast<unsigned char*>(&packed);
#pragma unroll
          for (int i = 0; i < 8; ++i) b[i] = __nv_cvt_float_to_fp8(v[i] * inv, __NV_SATFINITE, __NV_E4M3);
          *reinterpret_cast<uint2*>(dst + col) = packed;
          if ((lane & 15) == 0) sdst[col / 128] = scale;
        }
      }
    }
    if (lane == 0) {
      scatter_idx[p] = slot;
      symm_at(c, recv_src_pair, d)[slot] = c.rank * num_pairs + p;  // provenance (debug / weights lookup)
    }
  }
  // all ranks' tokens have landed when every CTA of every rank passed this point
  grid_barrier(grid_counter);
  if (blockIdx.x == 0) {
    for (int e = threadIdx.x; e < num_experts; e += blockDim.x) cursors[e] = 0;
    barrier_all_block(c, flags, base + 1);
  }
  phase_end(ps, base, 2);
}

// ------------------------------------------------------------------ EP combine
// expert_out (symmetric, rows in dispatch order) -> out[t] = sum_k w[t,k] * expert_out@rank(e_k)[scatter_idx[t,k]].
template <typename Tag, typename T>
__global__ void __launch_bounds__(512) ep_combine_kernel(SymmCtx c, const T* expert_out, const int* __restrict__ topk_ids,
                                                         const int* __restrict__ scatter_idx, const float* __restrict__ weights,
                                                         T* __restrict__ out, int tokens, int topk, int hidden, int num_experts,
                                                         uint32_t* flags, uint32_t* grid_counter, PhaseState ps) {
  const uint32_t base = phase_begin(ps);
  const int epr = num_experts / c.world;
  // every rank's expert outputs must be in place before anyone pulls
  if (blockIdx.x == 0) barrier_all_block(c, flags, base + 1);
  grid_barrier(grid_counter);
  const int row_vecs = hidden / 8;
  const size_t total = size_t(tokens) * row_vecs;
  for (size_t idx = blockIdx.x * size_t(blockDim.x) + threadIdx.x; idx < total; idx += size_t(gridDim.x) * blockDim.x) {
    const int t = int(idx / row_vecs), v = int(idx % row_vecs);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < topk; ++k) {
      const int e = topk_ids[t * topk + k];
      const int row = scatter_idx[t * topk + k];
      if (row < 0) continue;
      const int4 raw = ptx::ld_v4(reinterpret_cast<const int4*>(symm_at(c, expert_out, e / epr) + size_t(row) * hidden) + v);
      float part[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      Pack<Tag>::add(part, raw);
      const float w = weights ? weights[t * topk + k] : 1.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = fmaf(w, part[i], acc[i]);
    }
    reinterpret_cast<int4*>(out)[idx] = Pack<Tag>::cvt(acc);
  }
  // nobody may overwrite its expert_out buffer before all ranks finished pulling
  grid_barrier(grid_counter);
  if (blockIdx.x == 0) barrier_all_block(c, flags, base + 2);
  phase_end(ps, base, 2);
}

}  // namespace tdist

using namespace tdist;

extern "C" {

int tdist_a2a(const tdist_symm_ctx* ctx, const void* send, const int* splits_cumsum, const float* send_scale, void* recv_data,
              int* recv_splits, float* recv_scale, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_m, int row_bytes,
              int experts_per_rank, int ctas_per_peer, void* stream) {
  if (row_bytes % 16 != 0) {
    set_error("a2a: row bytes must be a multiple of 16");
    return -1;
  }
  const SymmCtx c = to_ctx(ctx);
  a2a_kernel<<<c.world * ctas_per_peer, 512, 0, static_cast<cudaStream_t>(stream)>>>(
      c, static_cast<const char*>(send), splits_cumsum, send_scale, static_cast<char*>(recv_data), recv_splits, recv_scale,
      flags, PhaseState{phase, done}, max_m, row_bytes, experts_per_rank, ctas_per_peer);
  return check_launch("a2a");
}

int tdist_ep_counts(const tdist_symm_ctx* ctx, const int* topk_ids, int num_pairs, int num_experts, int* full_counts,
                    uint32_t* flags, uint32_t* phase, uint32_t* done, void* stream) {
  const SymmCtx c = to_ctx(ctx);
  ep_counts_kernel<<<1, 1024, num_experts * sizeof(int), static_cast<cudaStream_t>(stream)>>>(
      c, topk_ids, num_pairs, num_experts, full_counts, flags, PhaseState{phase, done});
  return check_launch("ep_counts");
}

int tdist_ep_dispatch(const tdist_symm_ctx* ctx, int dtype, const void* x, const int* topk_ids, const int* full_counts,
                      void* recv_buf, float* recv_scale, int* recv_src_pair, int* scatter_idx, int* cursors,
                      int* recv_expert_counts, int* num_recv, int tokens, int topk, int hidden, int num_experts, int fp8,
                      uint32_t* flags, uint32_t* grid_counter, uint32_t* phase, uint32_t* done, int num_ctas, void* stream) {
  const SymmCtx c = to_ctx(ctx);
  if (hidden % 256 != 0 || num_experts % c.world != 0) {
    set_error("ep_dispatch: hidden must be a multiple of 256 and experts divisible by world size");
    return -1;
  }
  if (num_ctas < 1) num_ctas = 32;
  if (num_ctas > 132) num_ctas = 132;  // all CTAs must be co-resident (grid barrier)
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t smem = num_experts * sizeof(int);
  const PhaseState ps{phase, done};
  if (dtype == TDIST_BF16)
    ep_dispatch_kernel<__nv_bfloat16><<<num_ctas, 1024, smem, s>>>(
        c, static_cast<const __nv_bfloat16*>(x), topk_ids, full_counts, static_cast<char*>(recv_buf), recv_scale, recv_src_pair,
        scatter_idx, cursors, recv_expert_counts, num_recv, tokens, topk, hidden, num_experts, fp8, flags, grid_counter, ps);
  else if (dtype == TDIST_F16)
    ep_dispatch_kernel<__half><<<num_ctas, 1024, smem, s>>>(
        c, static_cast<const __half*>(x), topk_ids, full_counts, static_cast<char*>(recv_buf), recv_scale, recv_src_pair,
        scatter_idx, cursors, recv_expert_counts, num_recv, tokens, topk, hidden, num_experts, fp8, flags, grid_counter, ps);
  else {
    set_error("ep_dispatch: dtype must be bf16 or fp16");
    return -1;
  }
  return check_launch("ep_dispatch");
}

int tdist_ep_combine(const tdist_symm_ctx* ctx, int dtype, const void* expert_out, const int* topk_ids, const int* scatter_idx,
                     const float* weights, void* out, int tokens, int topk, int hidden, int num_experts, uint32_t* flags,
                     uint32_t* grid_counter, uint32_t* phase, uint32_t* done, int num_ctas, void* stream) {
  const SymmCtx c = to_ctx(ctx);
  if (hidden % 8 != 0) {
    set_error("ep_combine: hidden must be a multiple of 8");
    return -1;
  }
  if (num_ctas < 1) num_ctas = 64;
  if (num_ctas > 132) num_ctas = 132;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const PhaseState ps{phase, done};
  if (dtype == TDIST_BF16)
    ep_combine_kernel<BF16, __nv_bfloat16><<<num_ctas, 512, 0, s>>>(
        c, static_cast<const __nv_bfloat16*>(expert_out), topk_ids, scatter_idx, weights, static_cast<__nv_bfloat16*>(out), tokens,
        topk, hidden, num_experts, flags, grid_counter, ps);
  else if (dtype == TDIST_F16)
    ep_combine_kernel<F16, __half><<<num_ctas, 512, 0, s>>>(c, static_cast<const __half*>(expert_out), topk_ids, scatter_idx,
                                                           weights, static_cast<__half*>(out), tokens, topk, hidden, num_experts,
                                                           flags, grid_counter, ps);
  else {
    set_error("ep_combine: dtype must be bf16 or fp16");
    return -1;
  }
  return check_launch("ep_combine");
}

}  // extern "C"
