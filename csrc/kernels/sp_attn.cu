// sp_attn.cu — varlen causal GQA flash-attention forward over an all-gathered K/V buffer with
// sequence-parallel (optionally zig-zag) query sharding.
//
// Reference: kernels/nvidia/sp_ag_attention_intra_node.py:255-425 (kernel_consumer_flash_attn_forward) and
// :186-252 (block-offset remap by source rank). Each rank holds 1/W of every sequence's queries; K/V of all ranks
// are gathered into one buffer in GLOBAL token order, so the kernel is a standard tiled online-softmax attention
// whose causal mask uses the query's global position:
//   plain sharding   : local row i of a sequence with global length L  ->  position rank * L/W + i
//   zig-zag sharding : rank r holds chunks r and 2W-1-r of 2W equal chunks (load-balanced causal work,
//                      reference :360-371); row i < h -> r*h + i, else (2W-1-r)*h + (i - h), h = L / (2W).
//
// Tensor cores through warp-level MMA (nvcuda::wmma, HMMA on sm_100a): a 64-row query tile per CTA (4 warps x 16
// rows), K/V tiles of 64 keys staged in shared memory with cp.async, fp32 online softmax, fp32 output accumulator
// kept in shared memory so that the per-row rescale needs no knowledge of the MMA fragment layout.
// (This round's version: correct and tensor-core based; a tcgen05/TMEM pipeline is the next step for this op.)
#include <cuda_bf16.h>
#include <math_constants.h>
#include <mma.h>

#include "../runtime/driver.h"
#include "common.cuh"

namespace tdist {

using namespace nvcuda;

constexpr int kBM = 64, kBN = 64, kD = 128;

struct SpAttnParams {
  const __nv_bfloat16* q;   // [total_q_local, Hq, D]
  const __nv_bfloat16* k;   // [total_k_global, Hkv, D] gathered, global order
  const __nv_bfloat16* v;
  __nv_bfloat16* out;       // [total_q_local, Hq, D]
  const int* cu_q;          // [B + 1] local query offsets
  const int* cu_k;          // [B + 1] global key offsets
  int Hq, Hkv, rank, world;
  int causal, zigzag;
  float scale_log2;
};

__device__ __forceinline__ int global_q_pos(int i, int Lq, int L, int rank, int world, int zigzag) {
  if (!zigzag) return rank * (L / world) + i;
  const int h = L / (2 * world);
  return i < h ? rank * h + i : (2 * world - 1 - rank) * h + (i - h);
}

__device__ __forceinline__ void cp_async16(void* smem, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ptx::smem_u32(smem)), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// grid (q tiles, Hq, B), 128 threads
__global__ void __launch_bounds__(128) sp_attn_fwd_kernel(const SpAttnParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem_raw);             // [64][128]
  __nv_bfloat16* sK = sQ + kBM * kD;                                           // [64][128]
  __nv_bfloat16* sV = sK + kBN * kD;                                           // [64][128]
  float* sS = reinterpret_cast<float*>(sV + kBN * kD);                         // [4 warps][16][64]
  __nv_bfloat16* sP = reinterpret_cast<__nv_bfloat16*>(sS + 4 * 16 * kBN);     // [4][16][64]
  float* sO = reinterpret_cast<float*>(sP + 4 * 16 * kBN);                     // [4][16][128]
  float* sAlpha = sO + 4 * 16 * kD;                                            // [64] rescale of the running output
  const int b = blockIdx.z, h = blockIdx.y, qt = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q_beg = p.cu_q[b], Lq = p.cu_q[b + 1] - q_beg;
  const int k_beg = p.cu_k[b], L = p.cu_k[b + 1] - k_beg;
  const int q0 = qt * kBM;
  if (q0 >= Lq) return;
  const int kvh = h / (p.Hq / p.Hkv);

  // load the Q tile (rows beyond Lq are zero)
  for (int idx = threadIdx.x; idx < kBM * (kD / 8); idx += blockDim.x) {
    const int r = idx / (kD / 8), c = idx % (kD / 8);
    int4 v = make_int4(0, 0, 0, 0);
    if (q0 + r < Lq) v = *reinterpret_cast<const int4*>(p.q + (size_t(q_beg + q0 + r) * p.Hq + h) * kD + c * 8);
    *reinterpret_cast<int4*>(sQ + r * kD + c * 8) = v;
  }
  for (int idx = threadIdx.x; idx < 4 * 16 * kD; idx += blockDim.x) sO[idx] = 0.f;
  // per-row softmax state: row = warp * 16 + (lane / 2); two lanes share a row (each scans 32 of the 64 columns)
  const int my_row = warp * 16 + (lane >> 1);
  const int my_qi = q0 + my_row;
  const int my_pos = my_qi < Lq ? global_q_pos(my_qi, Lq, L, p.rank, p.world, p.zigzag) : -1;
  float m_i = -CUDART_INF_F, l_i = 0.f;
  // furthest key any row of this tile may see
  int kv_end = L;
  if (p.causal) {
    int hi = 0;
    for (int r = 0; r < kBM; r += kBM - 1) {  // positions are monotone inside each zig-zag half: check both ends + split
      if (q0 + r < Lq) hi = max(hi, global_q_pos(q0 + r, Lq, L, p.rank, p.world, p.zigzag));
    }
    const int last = min(q0 + kBM, Lq) - 1;
    hi = max(hi, global_q_pos(last, Lq, L, p.rank, p.world, p.zigzag));
    if (p.zigzag) {
      const int hh = L / (2 * p.world);
      if (q0 < hh && last >= hh) hi = max(hi, global_q_pos(hh - 1, Lq, L, p.rank, p.world, 1));
    }
    kv_end = min(L, hi + 1);
  }
  __syncthreads();

  for (int n0 = 0; n0 < kv_end; n0 += kBN) {
    // stage K and V tiles
    for (int idx = threadIdx.x; idx < kBN * (kD / 8); idx += blockDim.x) {
      const int r = idx / (kD / 8), c = idx % (kD / 8);
      const int kp = n0 + r;
      __nv_bfloat16* dk = sK + r * kD + c * 8;
      __nv_bfloat16* dv = sV + r * kD + c * 8;
      if (kp < L) {
        const size_t off = (size_t(k_beg + kp) * p.Hkv + kvh) * kD + c * 8;
        cp_async16(dk, p.k + off);
        cp_async16(dv, p.v + off);
      } else {
        *reinterpret_cast<int4*>(dk) = make_int4(0, 0, 0, 0);
        *reinterpret_cast<int4*>(dv) = make_int4(0, 0, 0, 0);
      }
    }
    cp_async_commit();
    cp_async_wait_all();
    __syncthreads();

    // S = Q_w K^T : 16 x 64 per warp
    float* S = sS + warp * 16 * kBN;
    {
      wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[4];
#pragma unroll
      for (int n = 0; n < 4; ++n) wmma::fill_fragment(acc[n], 0.f);
#pragma unroll
      for (int kk = 0; kk < kD / 16; ++kk) {
        wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> a;
        wmma::load_matrix_sync(a, sQ + warp * 16 * kD + kk * 16, kD);
#pragma unroll
        for (int n = 0; n < 4; ++n) {
          wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::col_major> bf;
          wmma::load_matrix_sync(bf, sK + n * 16 * kD + kk * 16, kD);  // K^T: element (k, n) = K[n][k]
          wmma::mma_sync(acc[n], a, bf, acc[n]);
        }
      }
#pragma unroll
      for (int n = 0; n < 4; ++n) wmma::store_matrix_sync(S + n * 16, acc[n], kBN, wmma::mem_row_major);
    }
    __syncwarp();
    // online softmax on my half row
    {
      const int r = lane >> 1, c0 = (lane & 1) * 32;
      float mx = -CUDART_INF_F;
      float vals[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int kp = n0 + c0 + j;
        const bool ok = my_pos >= 0 && kp < L && (!p.causal || kp <= my_pos);
        vals[j] = ok ? S[r * kBN + c0 + j] * p.scale_log2 : -CUDART_INF_F;
        mx = fmaxf(mx, vals[j]);
      }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      const float m_new = fmaxf(m_i, mx);
      const float alpha = m_new == -CUDART_INF_F ? 1.f : exp2f(m_i - m_new);
      float sum = 0.f;
      __nv_bfloat16* P = sP + warp * 16 * kBN;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float e = m_new == -CUDART_INF_F ? 0.f : exp2f(vals[j] - m_new);
        const __nv_bfloat16 eb = __float2bfloat16(e);
        P[r * kBN + c0 + j] = eb;
        sum += __bfloat162float(eb);  // the sum uses the rounded probabilities that multiply V
      }
      sum += __shfl_xor_sync(0xffffffffu, sum, 1);
      l_i = l_i * alpha + sum;
      m_i = m_new;
      if ((lane & 1) == 0) sAlpha[my_row] = alpha;
    }
    __syncwarp();
    // rescale the running output rows of this warp, then O += P V
    float* O = sO + warp * 16 * kD;
    for (int idx = lane; idx < 16 * kD; idx += 32) O[idx] *= sAlpha[warp * 16 + idx / kD];
    __syncwarp();
    {
      const __nv_bfloat16* P = sP + warp * 16 * kBN;
#pragma unroll
      for (int d = 0; d < kD / 16; ++d) {
        wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc;
        wmma::load_matrix_sync(acc, O + d * 16, kD, wmma::mem_row_major);
#pragma unroll
        for (int kk = 0; kk < kBN / 16; ++kk) {
          wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> a;
          wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> bf;
          wmma::load_matrix_sync(a, P + kk * 16, kBN);
          wmma::load_matrix_sync(bf, sV + kk * 16 * kD + d * 16, kD);
          wmma::mma_sync(acc, a, bf, acc);
        }
        wmma::store_matrix_sync(O + d * 16, acc, kD, wmma::mem_row_major);
      }
    }
    __syncthreads();  // all warps are done with sK / sV before the next tile overwrites them
  }
  // normalise and write
  if ((lane & 1) == 0) sAlpha[my_row] = l_i > 0.f ? 1.f / l_i : 0.f;
  __syncwarp();
  const float* O = sO + warp * 16 * kD;
  for (int idx = lane; idx < 16 * (kD / 8); idx += 32) {
    const int r = idx / (kD / 8), c = idx % (kD / 8);
    const int qi = q0 + warp * 16 + r;
    if (qi >= Lq) continue;
    const float inv = sAlpha[warp * 16 + r];
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = O[r * kD + c * 8 + j] * inv;
    int4 o;
    o.x = ptx::pack_bf16x2(f[0], f[1]);
    o.y = ptx::pack_bf16x2(f[2], f[3]);
    o.z = ptx::pack_bf16x2(f[4], f[5]);
    o.w = ptx::pack_bf16x2(f[6], f[7]);
    *reinterpret_cast<int4*>(p.out + (size_t(q_beg + qi) * p.Hq + h) * kD + c * 8) = o;
  }
}

// scatter this rank's K/V shard rows into GLOBAL order in the gathered buffer of `peer` (one launch per peer, or
// peer == rank for the local copy). plain: rows of sequence b go to k_beg + rank*L/W + i; zig-zag: two half chunks.
__global__ void __launch_bounds__(256) sp_kv_scatter_kernel(const int4* __restrict__ src, int4* dst, const int* cu_k_global,
                                                           int B, int row_vecs, int rank, int world, int zigzag) {
  const int b = blockIdx.y;
  const int k_beg = cu_k_global[b], L = cu_k_global[b + 1] - k_beg;
  const int Ll = L / world;
  const int src_beg = k_beg / world;  // every sequence is split evenly, so local offsets are global / W
  const size_t total = size_t(Ll) * row_vecs;
  for (size_t idx = blockIdx.x * size_t(blockDim.x) + threadIdx.x; idx < total; idx += size_t(gridDim.x) * blockDim.x) {
    const int i = int(idx / row_vecs), c = int(idx % row_vecs);
    const int gp = global_q_pos(i, Ll, L, rank, world, zigzag);
    ptx::st_v4(dst + size_t(k_beg + gp) * row_vecs + c, ptx::ld_nc_v4(src + size_t(src_beg + i) * row_vecs + c));
  }
}

}  // namespace tdist

using namespace tdist;

extern "C" {

int tdist_sp_attn_fwd(const void* q, const void* k_gathered, const void* v_gathered, void* out, const int* cu_q, const int* cu_k,
                      int B, int max_q_local, int Hq, int Hkv, int head_dim, int rank, int world, int causal, int zigzag,
                      float sm_scale, void* stream) {
  if (head_dim != kD || Hq % Hkv != 0) {
    set_error("sp_attn_fwd: head_dim must be 128 and Hq a multiple of Hkv");
    return -1;
  }
  SpAttnParams p{static_cast<const __nv_bfloat16*>(q), static_cast<const __nv_bfloat16*>(k_gathered),
                 static_cast<const __nv_bfloat16*>(v_gathered), static_cast<__nv_bfloat16*>(out), cu_q, cu_k, Hq, Hkv, rank, world,
                 causal, zigzag, sm_scale * 1.4426950408889634f};
  const size_t smem = size_t(kBM + 2 * kBN) * kD * 2 + 4 * 16 * kBN * 4 + 4 * 16 * kBN * 2 + 4 * 16 * kD * 4 + kBM * 4;
  static bool attr = false;
  if (!attr) {
    TDIST_RT_TRY(cudaFuncSetAttribute(sp_attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    attr = true;
  }
  dim3 grid((max_q_local + kBM - 1) / kBM, Hq, B);
  sp_attn_fwd_kernel<<<grid, 128, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("sp_attn_fwd");
}

int tdist_sp_kv_scatter(const void* src, void* dst, const int* cu_k_global, int B, int max_local_rows, int row_bytes, int rank,
                        int world, int zigzag, void* stream) {
  if (row_bytes % 16 != 0) {
    set_error("sp_kv_scatter: row bytes must be a multiple of 16");
    return -1;
  }
  const int row_vecs = row_bytes / 16;
  int gx = int((size_t(max_local_rows) * row_vecs + 255) / 256);
  if (gx > 256) gx = 256;
  if (gx < 1) gx = 1;
  sp_kv_scatter_kernel<<<dim3(gx, B), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const int4*>(src), static_cast<int4*>(dst), cu_k_global, B, row_vecs, rank, world, zigzag);
  return check_launch("sp_kv_scatter");
}

}  // extern "C"
