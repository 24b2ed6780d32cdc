// moe.cu — MoE index preparation, row gather, top-k weighted reduce and the fused
// top-k-reduce + reduce-scatter.
//
// Reference counterparts: csrc/lib/moe_utils.cu:60-312 (dead CUDA prototype of the align/sort),
// kernels/nvidia/allgather_group_gemm.py:84-165 (calc_sorted_gather_index_kernel),
// kernels/nvidia/moe_utils.py:50-405 (index + reduce_topk kernels),
// kernels/nvidia/moe_reduce_rs.py:292-376 (reduce_topk_reduce_scatter_a2a_intra_node_kernel).
//
// Layout convention: "pairs" are flattened (token, k) indices p = token * topk + k. After align/sort,
// row i of the expert-sorted (block-padded) matrix holds pair sorted_pair_ids[i] (or `num_pairs` for a
// padding row) and pair p sits at row pair_to_row[p].
#include "common.cuh"

namespace tdist {

// ------------------------------------------------------------------ align + sort (single CTA)
// counts -> per-expert padded offsets -> scatter. Order inside an expert is arbitrary (atomic cursor);
// every row is computed independently downstream, so results do not depend on it.
__global__ void __launch_bounds__(1024) moe_align_sort_kernel(const int* __restrict__ topk_ids, int num_pairs,
                                                              int num_experts, int block_m, int* sorted_pair_ids,
                                                              int* pair_to_row, int* block_expert, int* meta,
                                                              int max_rows) {
  extern __shared__ int sh[];  // counts[E] | offsets[E + 1] | cursor[E]
  int* counts = sh;
  int* offsets = sh + num_experts;
  int* cursor = offsets + num_experts + 1;
  for (int e = threadIdx.x; e < num_experts; e += blockDim.x) counts[e] = 0;
  __syncthreads();
  for (int p = threadIdx.x; p < num_pairs; p += blockDim.x) {
    const int e = topk_ids[p];
    if (e >= 0 && e < num_experts) atomicAdd(&counts[e], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int e = 0; e < num_experts; ++e) {
      offsets[e] = acc;
      acc += (counts[e] + block_m - 1) / block_m * block_m;
    }
    offsets[num_experts] = acc;
    meta[0] = acc;            // rows after padding
    meta[1] = acc / block_m;  // blocks
  }
  __syncthreads();
  const int total = offsets[num_experts];
  for (int e = threadIdx.x; e < num_experts; e += blockDim.x) {
    cursor[e] = offsets[e];
    for (int b = offsets[e] / block_m; b < offsets[e + 1] / block_m; ++b) block_expert[b] = e;
  }
  for (int b = total / block_m + threadIdx.x; b < max_rows / block_m; b += blockDim.x) block_expert[b] = -1;
  for (int i = threadIdx.x; i < max_rows; i += blockDim.x) sorted_pair_ids[i] = num_pairs;
  __syncthreads();
  for (int p = threadIdx.x; p < num_pairs; p += blockDim.x) {
    const int e = topk_ids[p];
    if (e >= 0 && e < num_experts) {
      const int row = atomicAdd(&cursor[e], 1);
      sorted_pair_ids[row] = p;
      pair_to_row[p] = row;
    } else {
      pair_to_row[p] = -1;
    }
  }
}

// ------------------------------------------------------------------ gather rows into expert-sorted order
// dst[i, :] = src[sorted_pair_ids[i] / topk, :] (zeros for padding rows). One warp per row.
__global__ void __launch_bounds__(256) moe_gather_rows_kernel(const int4* __restrict__ src, int4* __restrict__ dst,
                                                              const int* __restrict__ sorted_pair_ids,
                                                              const int* __restrict__ meta, int num_pairs, int topk,
                                                              int row_vecs) {
  const int rows = meta[0];
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int i = warp; i < rows; i += nwarps) {
    const int p = sorted_pair_ids[i];
    int4* d = dst + size_t(i) * row_vecs;
    if (p >= num_pairs) {
      for (int v = lane; v < row_vecs; v += 32) d[v] = make_int4(0, 0, 0, 0);
    } else {
      const int4* s = src + size_t(p / topk) * row_vecs;
      for (int v = lane; v < row_vecs; v += 32) d[v] = ptx::ld_nc_v4(s + v);
    }
  }
}

// Stage-filtered form for the stage-gated AG + grouped GEMM (allgather_group_gemm.py, overlap mode): only the rows of
// blocks whose LAST needed shard arrives at `stage` (block_stage[i / block_m] == stage) are written; all other rows of
// dst are left untouched, so consecutive stages fill disjoint row blocks of the same buffer.
__global__ void __launch_bounds__(256) moe_gather_rows_stage_kernel(const int4* __restrict__ src, int4* __restrict__ dst,
                                                                    const int* __restrict__ sorted_pair_ids,
                                                                    const int* __restrict__ meta,
                                                                    const int* __restrict__ block_stage, int stage, int block_m,
                                                                    int num_pairs, int topk, int row_vecs) {
  const int rows = meta[0];
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int i = warp; i < rows; i += nwarps) {
    if (block_stage[i / block_m] != stage) continue;
    const int p = sorted_pair_ids[i];
    int4* d = dst + size_t(i) * row_vecs;
    if (p >= num_pairs) {
      for (int v = lane; v < row_vecs; v += 32) d[v] = make_int4(0, 0, 0, 0);
    } else {
      const int4* s = src + size_t(p / topk) * row_vecs;
      for (int v = lane; v < row_vecs; v += 32) d[v] = ptx::ld_nc_v4(s + v);
    }
  }
}

// ------------------------------------------------------------------ top-k weighted reduce
// out[t, :] = sum_k w[t, k] * y[pair_to_row[t * topk + k], :]. One warp per token.
template <typename Tag>
__device__ __forceinline__ void topk_reduce_row(const int4* __restrict__ y, const int* __restrict__ pair_to_row,
                                                const float* __restrict__ weights, int t, int topk, int row_vecs,
                                                int v, float (&acc)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int k = 0; k < topk; ++k) {
    const int row = pair_to_row[t * topk + k];
    if (row < 0) continue;
    float part[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    Pack<Tag>::add(part, ptx::ld_nc_v4(y + size_t(row) * row_vecs + v));
    const float w = weights ? weights[t * topk + k] : 1.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = fmaf(w, part[i], acc[i]);
  }
}

template <typename Tag>
__global__ void __launch_bounds__(256) moe_topk_reduce_kernel(const int4* __restrict__ y, int4* __restrict__ out,
                                                              const int* __restrict__ pair_to_row,
                                                              const float* __restrict__ weights, int tokens, int topk,
                                                              int row_vecs) {
  const size_t total = size_t(tokens) * row_vecs;
  for (size_t idx = blockIdx.x * size_t(blockDim.x) + threadIdx.x; idx < total; idx += size_t(gridDim.x) * blockDim.x) {
    const int t = int(idx / row_vecs), v = int(idx % row_vecs);
    float acc[8];
    topk_reduce_row<Tag>(y, pair_to_row, weights, t, topk, row_vecs, v, acc);
    out[idx] = Pack<Tag>::cvt(acc);
  }
}

// ------------------------------------------------------------------ fused top-k reduce + reduce-scatter
// Every rank holds y (expert-sorted partial outputs of ALL tokens, FFN dim sharded). Token segment s
// (tokens [s*T/W, (s+1)*T/W)) belongs to rank s. Each CTA reduces its slice of segment s over top-k and
// stores it into slot `rank` of rank s's scatter buffer; after the per-CTA flag exchange the owner sums
// the W slots. scat (symm): [2 parity][W][seg_tokens * row_vecs].
template <typename Tag>
__global__ void __launch_bounds__(512) moe_reduce_rs_kernel(SymmCtx c, const int4* __restrict__ y, int4* scat,
                                                            int4* __restrict__ out, const int* __restrict__ pair_to_row,
                                                            const float* __restrict__ weights, int tokens, int topk,
                                                            int row_vecs, size_t parity_vecs, uint32_t* flags, PhaseState ps) {
  const uint32_t base = phase_begin(ps);
  const int W = c.world;
  const int seg_tokens = tokens / W;
  const size_t seg_vecs = size_t(seg_tokens) * row_vecs;
  int4* buf = scat + ((base >> 1) & 1) * parity_vecs;  // fixed distance between the parity halves of the workspace
  const size_t per_cta = (seg_vecs + gridDim.x - 1) / gridDim.x;
  const size_t beg = blockIdx.x * per_cta;
  const size_t end = min(beg + per_cta, seg_vecs);
#pragma unroll 1
  for (int p = 1; p <= W; ++p) {
    const int o = (c.rank + p) % W;  // own segment last
    int4* dst = symm_at(c, buf, o) + c.rank * seg_vecs;
    for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
      const int t = o * seg_tokens + int(i / row_vecs), v = int(i % row_vecs);
      float acc[8];
      topk_reduce_row<Tag>(y, pair_to_row, weights, t, topk, row_vecs, v, acc);
      ptx::st_v4(dst + i, Pack<Tag>::cvt(acc));
    }
  }
  barrier_all_block(c, flags + blockIdx.x * TDIST_MAX_RANKS, base + 1);
  for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int s = 0; s < W; ++s) Pack<Tag>::add(acc, ptx::ld_v4(buf + s * seg_vecs + i));
    out[i] = Pack<Tag>::cvt(acc);
  }
  phase_end(ps, base, 2);
}

}  // namespace tdist

using namespace tdist;

// ------------------------------------------------------------------ AG+MoE tile order (reference K45)
// For every 128-row block of the expert-sorted layout: which source ranks' tokens it needs ([seg_start, seg_end] in
// arrival order, where rank r's own shard arrives first, then r+1, r+2, ...), and a stable order of the blocks by the
// stage at which their LAST needed shard arrives. A grouped GEMM that walks the blocks in this order can start on
// the blocks fed by early shards while later shards are still in flight (threadblock_swizzle_ag_moe_triton.py:95-216).
// One CTA; blocks and ranks are small numbers (<= a few thousand blocks, <= 16 ranks).
__global__ void __launch_bounds__(1024) moe_tile_swizzle_kernel(const int* __restrict__ sorted_pair_ids,
                                                                const int* __restrict__ meta, int num_pairs, int topk,
                                                                int tokens_per_rank, int rank, int world, int block_m,
                                                                int* __restrict__ block_order, int* __restrict__ seg_start,
                                                                int* __restrict__ seg_end, int* __restrict__ stage_count,
                                                                int max_blocks) {
  __shared__ int s_count[TDIST_MAX_RANKS + 1];
  __shared__ int s_offset[TDIST_MAX_RANKS + 1];
  const int nblk = min(meta[1], max_blocks);
  if (threadIdx.x <= TDIST_MAX_RANKS) s_count[threadIdx.x] = 0;
  __syncthreads();
  // stage of a block = max over its rows of ((src_rank - rank) mod world); empty / padding-only blocks get stage 0
  for (int b = threadIdx.x; b < nblk; b += blockDim.x) {
    int lo = world, hi = -1;
    for (int r = 0; r < block_m; ++r) {
      const int pid = sorted_pair_ids[b * block_m + r];
      if (pid < num_pairs) {
        const int src = (pid / topk) / tokens_per_rank;
        const int st = (src - rank + world) % world;
        lo = min(lo, st);
        hi = max(hi, st);
      }
    }
    if (hi < 0) lo = hi = 0;
    seg_start[b] = lo;
    seg_end[b] = hi;
    atomicAdd(&s_count[hi], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int s = 0; s < world; ++s) {
      s_offset[s] = run;
      stage_count[s] = s_count[s];
      run += s_count[s];
    }
  }
  __syncthreads();
  // stable counting sort by stage: block b goes behind all blocks of the same stage with a smaller index
  for (int b = threadIdx.x; b < nblk; b += blockDim.x) {
    const int st = seg_end[b];
    int before = 0;
    for (int j = 0; j < b; ++j) before += seg_end[j] == st;
    block_order[s_offset[st] + before] = b;
  }
}

extern "C" {

int tdist_moe_tile_swizzle(const int* sorted_pair_ids, const int* meta, int num_pairs, int topk, int tokens_per_rank, int rank,
                           int world, int block_m, int* block_order, int* seg_start, int* seg_end, int* stage_count,
                           int max_blocks, void* stream) {
  if (world < 1 || world > TDIST_MAX_RANKS || tokens_per_rank < 1 || topk < 1) {
    tdist::set_error("moe_tile_swizzle: bad arguments");
    return -1;
  }
  moe_tile_swizzle_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(
      sorted_pair_ids, meta, num_pairs, topk, tokens_per_rank, rank, world, block_m, block_order, seg_start, seg_end, stage_count,
      max_blocks);
  return tdist::check_launch("moe_tile_swizzle");
}

int tdist_moe_align_sort(const int* topk_ids, int num_pairs, int num_experts, int block_m, int* sorted_pair_ids,
                         int* pair_to_row, int* block_expert, int* meta, int max_rows, void* stream) {
  if (max_rows % block_m != 0) {
    set_error("moe_align_sort: max_rows must be a multiple of block_m");
    return -1;
  }
  const size_t smem = (size_t(num_experts) * 3 + 1) * sizeof(int);
  if (smem > 48 * 1024) {
    set_error("moe_align_sort: too many experts");
    return -1;
  }
  moe_align_sort_kernel<<<1, 1024, smem, static_cast<cudaStream_t>(stream)>>>(
      topk_ids, num_pairs, num_experts, block_m, sorted_pair_ids, pair_to_row, block_expert, meta, max_rows);
  return check_launch("moe_align_sort");
}

int tdist_moe_gather_rows(const void* src, void* dst, const int* sorted_pair_ids, const int* meta, int num_pairs,
                          int topk, int row_bytes, int max_rows, void* stream) {
  if (row_bytes % 16 != 0) {
    set_error("moe_gather_rows: row bytes must be a multiple of 16");
    return -1;
  }
  int grid = (max_rows + 7) / 8;
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  moe_gather_rows_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const int4*>(src), static_cast<int4*>(dst), sorted_pair_ids, meta, num_pairs, topk, row_bytes / 16);
  return check_launch("moe_gather_rows");
}

int tdist_moe_gather_rows_stage(const void* src, void* dst, const int* sorted_pair_ids, const int* meta, const int* block_stage,
                                int stage, int block_m, int num_pairs, int topk, int row_bytes, int max_rows, void* stream) {
  if (row_bytes % 16 != 0 || block_m < 1) {
    set_error("moe_gather_rows_stage: row bytes must be a multiple of 16");
    return -1;
  }
  int grid = (max_rows + 7) / 8;
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  moe_gather_rows_stage_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const int4*>(src), static_cast<int4*>(dst), sorted_pair_ids, meta, block_stage, stage, block_m, num_pairs, topk,
      row_bytes / 16);
  return check_launch("moe_gather_rows_stage");
}

int tdist_moe_topk_reduce(int dtype, const void* y, void* out, const int* pair_to_row, const float* weights, int tokens,
                          int topk, int row_bytes, void* stream) {
  if (row_bytes % 16 != 0) {
    set_error("moe_topk_reduce: row bytes must be a multiple of 16");
    return -1;
  }
  const int row_vecs = row_bytes / 16;
  size_t total = size_t(tokens) * row_vecs;
  int grid = int((total + 255) / 256);
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  TDIST_DISPATCH_DTYPE(dtype, {
    moe_topk_reduce_kernel<Tag><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const int4*>(y), static_cast<int4*>(out), pair_to_row, weights, tokens, topk, row_vecs);
  });
  return check_launch("moe_topk_reduce");
}

int tdist_moe_reduce_rs(const tdist_symm_ctx* ctx, int dtype, const void* y, void* scatter_ws, size_t ws_bytes, void* out,
                        const int* pair_to_row, const float* weights, int tokens, int topk, int row_bytes, uint32_t* flags,
                        uint32_t* phase, uint32_t* done, int max_ctas, void* stream) {
  const SymmCtx c = to_ctx(ctx);
  if (row_bytes % 16 != 0 || tokens % c.world != 0) {
    set_error("moe_reduce_rs: row bytes must be a multiple of 16 and tokens divisible by world size");
    return -1;
  }
  const int row_vecs = row_bytes / 16;
  const size_t seg_vecs = size_t(tokens / c.world) * row_vecs;
  const size_t parity_vecs = ws_bytes / 16 / 2;
  if (parity_vecs < size_t(c.world) * seg_vecs) {
    set_error("moe_reduce_rs: workspace too small (need 2 * tokens * row_bytes)");
    return -1;
  }
  if (max_ctas > TDIST_MAX_CTAS) max_ctas = TDIST_MAX_CTAS;
  int grid = int((seg_vecs + 511) / 512);
  if (grid > max_ctas) grid = max_ctas;
  if (grid < 1) grid = 1;
  TDIST_DISPATCH_DTYPE(dtype, {
    moe_reduce_rs_kernel<Tag><<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(
        c, static_cast<const int4*>(y), static_cast<int4*>(scatter_ws), static_cast<int4*>(out), pair_to_row, weights,
        tokens, topk, row_vecs, parity_vecs, flags, PhaseState{phase, done});
  });
  return check_launch("moe_reduce_rs");
}

}  // extern "C"
