// sp_attn_sm100.cu — tcgen05 / TMEM / TMA flash-attention forward (head_dim 128, bf16): varlen, GQA, causal, with the
// sequence-parallel query-position mapping (plain or zig-zag) of sp_attn.cu. Used by the SP all-gather attention
// (reference kernels/nvidia/sp_ag_attention_intra_node.py:255-425, which is tl.dot based) and by TP_Attn's prefill.
//
// One CTA = one 128-row query tile of one (sequence, head); K/V tiles of 128 keys.
//   warp 8    TMA producer: Q once, then K_j and V_j into two-stage rings of their own (K_j is released right after
//             S_j = Q K_j^T, long before V_j, so K_{j+2} is already loading while tile j's softmax runs)
//   warp 9    allocates TMEM and issues the UMMAs: S_{j+1} = Q K_{j+1}^T is issued BEFORE P_j V_j, so the tensor core
//             computes the next scores while the softmax warps are busy with the current ones
//   warps 0-7 softmax: TWO threads per query row (warps w and w + 4 share TMEM lanes), each with 64 keys of the tile and
//             64 output dims in registers. They read S_j from TMEM once, exchange the row maximum through shared
//             memory, exponentiate and write P_j (bf16) into shared memory in the K-major 128B-swizzled operand
//             layout; the partial product P_j V_j comes back through TMEM and is folded into the fp32 output
//             registers one tile later (O = O * alpha + PV), i.e. the rescale never touches TMEM.
//             (One thread per row with 128 scores + 128 outputs spilled and ran at 200 TFLOP/s.)
// V is consumed as an MN-major B operand straight from its [keys][d] layout (no transpose anywhere).
// TMEM: S double buffered (2 x 128 columns) + PV (128 columns).
#include <cuda_bf16.h>
#include <math_constants.h>

#include "../runtime/driver.h"
#include "common.cuh"

namespace tdist {

constexpr int kAtBM = 128, kAtBN = 128, kAtD = 128;
constexpr int kAtThreads = 320;
constexpr int kAtPanel = 128 * 128;  // bytes of one [128 rows][64 x bf16] swizzled panel

struct AttnTcParams {
  CUtensorMap tma_q, tma_k, tma_v;  // [rows, H * 128] bf16, box {64, 128}, 128B swizzle
  __nv_bfloat16* out;               // [total_q_local, Hq, D]
  const int* cu_q;                  // [B + 1] local query offsets
  const int* cu_k;                  // [B + 1] global key offsets
  int Hq, Hkv, rank, world;
  int causal, zigzag;
  int k_batch_stride;  // > 0: sequence b's keys start at row b * k_batch_stride (dense cache); 0: at cu_k[b]
  float scale_log2;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ int attn_q_pos(int i, int Lq, int L, int rank, int world, int zigzag) {
  if (world == 1) return i + (L - Lq);  // one rank: the queries are the LAST Lq positions of the sequence (prefill / chunked prefill)
  if (!zigzag) return rank * (L / world) + i;
  const int h = L / (2 * world);
  return i < h ? rank * h + i : (2 * world - 1 - rank) * h + (i - h);
}

// shared-memory matrix descriptor, 128B swizzle, explicit leading / stride byte offsets
__device__ __forceinline__ uint64_t attn_desc(const void* smem, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((ptx::smem_u32(smem) & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;
  d |= static_cast<uint64_t>(sbo_bytes >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

struct AttnSmem {
  static constexpr int kQ = 0;                       // 2 panels
  static constexpr int kK = kQ + 2 * kAtPanel;       // 2 stages x 2 panels
  static constexpr int kV = kK + 4 * kAtPanel;       // 2 stages x 2 panels
  static constexpr int kP = kV + 4 * kAtPanel;       // 2 panels
  static constexpr int kBar = kP + 2 * kAtPanel;     // barriers
  static constexpr int kTotal = kBar + 24 * 8 + 2 * 128 * 4 + 16;  // barriers, TMEM pointer, row-exchange floats
};

__global__ void __launch_bounds__(kAtThreads, 1) attn_fwd_tc_kernel(const __grid_constant__ AttnTcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y;
  const int qt = gridDim.x - 1 - blockIdx.x;  // causal: the last query tiles see the most keys, start them first
  const int q_beg = p.cu_q[b], Lq = p.cu_q[b + 1] - q_beg;
  const int L = p.cu_k[b + 1] - p.cu_k[b];
  const int k_beg = p.k_batch_stride > 0 ? b * p.k_batch_stride : p.cu_k[b];  // dense cache [B][S] or packed varlen
  const int q0 = qt * kAtBM;
  if (q0 >= Lq) return;  // uniform for the CTA: nothing allocated yet
  const int kvh = h / (p.Hq / p.Hkv);

  uint8_t* sQ = smem + AttnSmem::kQ;
  uint8_t* sK = smem + AttnSmem::kK;
  uint8_t* sV = smem + AttnSmem::kV;
  uint8_t* sP = smem + AttnSmem::kP;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnSmem::kBar);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;   // [2]
  uint64_t* k_empty = bars + 3;  // [2]
  uint64_t* v_full = bars + 5;   // [2]
  uint64_t* v_empty = bars + 7;  // [2]
  uint64_t* s_full = bars + 9;   // [2]
  uint64_t* s_empty = bars + 11; // [2]
  uint64_t* p_full = bars + 13;
  uint64_t* p_empty = bars + 14;
  uint64_t* pv_full = bars + 15;
  uint64_t* pv_empty = bars + 16;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 20);

  // furthest key any row of this tile may see (same rule as sp_attn.cu)
  int kv_end = L;
  if (p.causal) {
    const int last = min(q0 + kAtBM, Lq) - 1;
    int hi = max(attn_q_pos(q0, Lq, L, p.rank, p.world, p.zigzag), attn_q_pos(last, Lq, L, p.rank, p.world, p.zigzag));
    if (p.zigzag) {
      const int hh = L / (2 * p.world);
      if (q0 < hh && last >= hh) hi = max(hi, attn_q_pos(hh - 1, Lq, L, p.rank, p.world, 1));
    }
    kv_end = min(L, hi + 1);
  }
  const int n_tiles = (kv_end + kAtBN - 1) / kAtBN;

  if (threadIdx.x == 0) {
    ptx::mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(k_full + i, 1);
      ptx::mbar_init(k_empty + i, 1);
      ptx::mbar_init(v_full + i, 1);
      ptx::mbar_init(v_empty + i, 1);
      ptx::mbar_init(s_full + i, 1);
      ptx::mbar_init(s_empty + i, 8);
    }
    ptx::mbar_init(p_full, 8);
    ptx::mbar_init(p_empty, 1);
    ptx::mbar_init(pv_full, 1);
    ptx::mbar_init(pv_empty, 8);
    ptx::fence_barrier_init();
  }
  if (warp == 9) ptx::tmem_alloc<512>(tmem_ptr_smem);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  constexpr uint32_t kTmemPV = 256;

  if (warp == 8) {
    // ======================================================== TMA producer
    if (lane == 0 && n_tiles > 0) {
      ptx::prefetch_tensormap(&p.tma_q);
      ptx::prefetch_tensormap(&p.tma_k);
      ptx::prefetch_tensormap(&p.tma_v);
      ptx::mbar_arrive_expect_tx(q_full, 2 * kAtPanel);
      ptx::tma_load_2d(sQ, &p.tma_q, q_full, h * kAtD, q_beg + q0);
      ptx::tma_load_2d(sQ + kAtPanel, &p.tma_q, q_full, h * kAtD + 64, q_beg + q0);
      auto load_k = [&](int j) {
        const int s = j & 1;
        ptx::mbar_wait(k_empty + s, ((j >> 1) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(k_full + s, 2 * kAtPanel);
        ptx::tma_load_2d(sK + s * 2 * kAtPanel, &p.tma_k, k_full + s, kvh * kAtD, k_beg + j * kAtBN);
        ptx::tma_load_2d(sK + s * 2 * kAtPanel + kAtPanel, &p.tma_k, k_full + s, kvh * kAtD + 64, k_beg + j * kAtBN);
      };
      load_k(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) load_k(j + 1);  // K runs one tile ahead of V: its buffer is released much earlier
        const int s = j & 1;
        ptx::mbar_wait(v_empty + s, ((j >> 1) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(v_full + s, 2 * kAtPanel);
        ptx::tma_load_2d(sV + s * 2 * kAtPanel, &p.tma_v, v_full + s, kvh * kAtD, k_beg + j * kAtBN);
        ptx::tma_load_2d(sV + s * 2 * kAtPanel + kAtPanel, &p.tma_v, v_full + s, kvh * kAtD + 64, k_beg + j * kAtBN);
      }
    }
  } else if (warp == 9) {
    // ======================================================== MMA issuer
    if (lane == 0 && n_tiles > 0) {
      const uint32_t idesc_qk = ptx::make_idesc_f16(kAtBM, kAtBN, true);
      const uint32_t idesc_pv = ptx::make_idesc_f16(kAtBM, kAtD, true) | (1u << 16);  // B (= V) is MN-major
      auto issue_qk = [&](int j) {
        const int s = j & 1;
        ptx::mbar_wait(k_full + s, (j >> 1) & 1);
        ptx::mbar_wait(s_empty + s, ((j >> 1) & 1) ^ 1);
        ptx::tcgen05_fence_after();
#pragma unroll
        for (int kp = 0; kp < 2; ++kp) {
          const uint64_t da = ptx::make_smem_desc_sw128(sQ + kp * kAtPanel);
          const uint64_t db = ptx::make_smem_desc_sw128(sK + s * 2 * kAtPanel + kp * kAtPanel);
#pragma unroll
          for (int k = 0; k < 4; ++k) ptx::umma_f16_ss(tmem_base + s * kAtBN, da + 2 * k, db + 2 * k, idesc_qk, (kp | k) != 0);
        }
        ptx::umma_commit(k_empty + s);
        ptx::umma_commit(s_full + s);
      };
      ptx::mbar_wait(q_full, 0);
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_qk(j + 1);  // next scores first: they are computed while the softmax of tile j runs
        const int s = j & 1;
        ptx::mbar_wait(p_full, j & 1);
        ptx::mbar_wait(v_full + s, (j >> 1) & 1);
        ptx::mbar_wait(pv_empty, (j & 1) ^ 1);
        ptx::tcgen05_fence_after();
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          // A = P: K-major panels of 64 keys; B = V[keys][d]: MN-major, 64-wide d groups 16 KB apart, 8-key groups 1 KB
          const uint64_t da = ptx::make_smem_desc_sw128(sP + (kk >> 2) * kAtPanel) + 2 * (kk & 3);
          const uint64_t db = attn_desc(sV + s * 2 * kAtPanel + kk * 2048, kAtPanel, 1024);
          ptx::umma_f16_ss(tmem_base + kTmemPV, da, db, idesc_pv, kk != 0);
        }
        ptx::umma_commit(v_empty + s);
        ptx::umma_commit(p_empty);
        ptx::umma_commit(pv_full);
      }
    }
  } else {
    // ======================================================== softmax + output: two threads per query row
    // Warps w and w + 4 own the same 32 TMEM lanes (= query rows); warp group 0 (w < 4) handles keys 0..63 of every
    // K/V tile and output dims 0..63, group 1 the other halves. The two threads of a row exchange their partial row
    // maximum through shared memory (one 64-thread named barrier per tile); the row sums are only combined at the end.
    const int half = warp >> 2;               // which 64 keys / 64 output dims
    const int quarter = warp & 3;             // TMEM lane quarter
    const int row = quarter * 32 + lane;      // query row of this thread
    const int my_qi = q0 + row;
    const int my_pos = my_qi < Lq ? attn_q_pos(my_qi, Lq, L, p.rank, p.world, p.zigzag) : -1;
    const int vis_end = p.causal ? min(L, my_pos + 1) : L;  // keys [0, vis_end) are visible to this row
    const uint32_t lane_base = tmem_base + (uint32_t(quarter * 32) << 16);
    float* xch = reinterpret_cast<float*>(bars + 21);  // [2 halves][128 rows] partial maxima, later partial sums
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
    float m_i = -CUDART_INF_F, l_i = 0.f, alpha_prev = 1.f;
    auto pair_sync = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(1 + quarter) : "memory"); };
    auto fold_pv = [&](int j) {  // O = O * alpha + P_j V_j (my 64 output dims)
      ptx::mbar_wait(pv_full, j & 1);
      ptx::tcgen05_fence_after();
      uint32_t v0[32], v1[32];
      ptx::tmem_ld_32x32b_x32(lane_base + kTmemPV + half * 64, v0);
      ptx::tmem_ld_32x32b_x32(lane_base + kTmemPV + half * 64 + 32, v1);
      ptx::tmem_ld_wait();
      ptx::tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(pv_empty);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        o[i] = fmaf(o[i], alpha_prev, __uint_as_float(v0[i]));
        o[32 + i] = fmaf(o[32 + i], alpha_prev, __uint_as_float(v1[i]));
      }
    };
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j & 1;
      const int k0 = j * kAtBN + half * 64;  // first key of my half tile
      ptx::mbar_wait(s_full + s, (j >> 1) & 1);
      ptx::tcgen05_fence_after();
      uint32_t sv[64];
      {
        uint32_t (&a0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sv[0]);
        uint32_t (&a1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sv[32]);
        ptx::tmem_ld_32x32b_x32(lane_base + s * kAtBN + half * 64, a0);
        ptx::tmem_ld_32x32b_x32(lane_base + s * kAtBN + half * 64 + 32, a1);
        ptx::tmem_ld_wait();
      }
      ptx::tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(s_empty + s);  // the scores are in registers: S buffer free for Q K_{j+2}^T
      const bool full = k0 + 64 <= vis_end;  // no masking needed in my half tile
      float mx = -CUDART_INF_F;
      if (full) {
        float m4[4] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};  // four independent chains
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
#pragma unroll
          for (int u = 0; u < 4; ++u) m4[u] = fmaxf(m4[u], __uint_as_float(sv[i + u]));
        }
        mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      } else {
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (k0 + i < vis_end) mx = fmaxf(mx, __uint_as_float(sv[i]));
      }
      xch[half * 128 + row] = mx;
      pair_sync();
      mx = fmaxf(mx, xch[(half ^ 1) * 128 + row]);
      const float m_new = fmaxf(m_i, mx * p.scale_log2);
      const float alpha = m_new == -CUDART_INF_F ? 1.f : exp2f(m_i - m_new);
      const float m_use = m_new == -CUDART_INF_F ? 0.f : m_new;
      // P = exp2(S * scale - m) -> bf16, kept in registers until the tensor core has finished reading the previous P
      // (the exponentials are the long part of a tile: they must not wait for P_{j-1} V_{j-1})
      float l4[4] = {0.f, 0.f, 0.f, 0.f};
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float p0 = ex2_approx(fmaf(__uint_as_float(sv[2 * i]), p.scale_log2, -m_use));
        float p1 = ex2_approx(fmaf(__uint_as_float(sv[2 * i + 1]), p.scale_log2, -m_use));
        if (!full) {
          if (k0 + 2 * i >= vis_end) p0 = 0.f;
          if (k0 + 2 * i + 1 >= vis_end) p1 = 0.f;
        }
        l4[i & 3] += p0 + p1;
        pk[i] = ptx::pack_bf16x2(p0, p1);
      }
      const float l_add = (l4[0] + l4[1]) + (l4[2] + l4[3]);
      if (j > 0) ptx::mbar_wait(p_empty, (j - 1) & 1);  // P_{j-1} V_{j-1} has finished reading the P buffer
      uint8_t* prow = sP + half * kAtPanel + row * 128;  // K-major panel `half`, 128B swizzle
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<int4*>(prow + ((c ^ (row & 7)) << 4)) = make_int4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      if (j * kAtBN + kAtBN > L) {
        // keys beyond the sequence get probability 0, but 0 * (whatever lies behind the sequence in the buffer) must
        // not produce NaN: clear those V rows of this tile (each thread: its key row, its 64 dims)
        ptx::mbar_wait(v_full + s, (j >> 1) & 1);
        if (j * kAtBN + row >= L) {
          int4* vr = reinterpret_cast<int4*>(sV + s * 2 * kAtPanel + half * kAtPanel + row * 128);
#pragma unroll
          for (int i = 0; i < 8; ++i) vr[i] = make_int4(0, 0, 0, 0);
        }
      }
      ptx::fence_proxy_async_shared();  // generic-proxy writes to shared memory (P, cleared V rows) -> visible to the tensor core
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
      l_i = l_i * alpha + l_add;
      m_i = m_new;
      if (j > 0) fold_pv(j - 1);  // the previous tile's P V, computed by the tensor core during this tile's softmax
      alpha_prev = alpha;
      pair_sync();  // both halves have read the exchanged maxima before the next tile overwrites them
    }
    if (n_tiles > 0) fold_pv(n_tiles - 1);
    xch[half * 128 + row] = l_i;
    pair_sync();
    const float l_tot = l_i + xch[(half ^ 1) * 128 + row];
    if (my_qi < Lq) {
      const float inv = l_tot > 0.f ? 1.f / l_tot : 0.f;
      __nv_bfloat16* orow = p.out + (size_t(q_beg + my_qi) * p.Hq + h) * kAtD + half * 64;
#pragma unroll
      for (int c = 0; c < 64; c += 8) {
        int4 w;
        w.x = ptx::pack_bf16x2(o[c] * inv, o[c + 1] * inv);
        w.y = ptx::pack_bf16x2(o[c + 2] * inv, o[c + 3] * inv);
        w.z = ptx::pack_bf16x2(o[c + 4] * inv, o[c + 5] * inv);
        w.w = ptx::pack_bf16x2(o[c + 6] * inv, o[c + 7] * inv);
        *reinterpret_cast<int4*>(orow + c) = w;
      }
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  if (warp == 9) ptx::tmem_dealloc<512>(tmem_base);
}

}  // namespace tdist

namespace tdist {
// [rows, cols] bf16 row-major (row stride ld elements), box = 64 columns x 128 rows, 128B swizzle
static int attn_tmap(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld) {
  const DriverApi& d = driver();
  if (!d.ok) {
    set_error("CUDA driver not available: " + d.err);
    return -1;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = d.cuTensorMapEncodeTiled(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (attention) failed: " + cu_err(r));
    return -1;
  }
  return 0;
}
}  // namespace tdist

using namespace tdist;

extern "C" int tdist_attn_fwd_tc(const void* q, const void* k, const void* v, void* out, const int* cu_q, const int* cu_k, int B,
                                 int max_q_local, int total_q, int total_k, int Hq, int Hkv, int head_dim, long long q_row_stride,
                                 long long kv_row_stride, int k_batch_stride, int rank, int world, int causal, int zigzag,
                                 float sm_scale, void* stream) {
  if (head_dim != kAtD || Hkv <= 0 || Hq % Hkv != 0) {
    set_error("attn_fwd_tc: head_dim must be 128 and Hq a multiple of Hkv");
    return -1;
  }
  if (B <= 0 || max_q_local <= 0 || total_q <= 0 || total_k <= 0) return 0;  // nothing to do
  if (q_row_stride % 8 != 0 || kv_row_stride % 8 != 0 || q_row_stride < (long long)Hq * kAtD || kv_row_stride < (long long)Hkv * kAtD) {
    set_error("attn_fwd_tc: row strides must cover all heads and be multiples of 8 elements");
    return -1;
  }
  AttnTcParams p;
  memset(&p, 0, sizeof(p));
  if (attn_tmap(&p.tma_q, q, total_q, uint64_t(Hq) * kAtD, q_row_stride) != 0) return -1;
  if (attn_tmap(&p.tma_k, k, total_k, uint64_t(Hkv) * kAtD, kv_row_stride) != 0) return -1;
  if (attn_tmap(&p.tma_v, v, total_k, uint64_t(Hkv) * kAtD, kv_row_stride) != 0) return -1;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.cu_q = cu_q;
  p.cu_k = cu_k;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.rank = rank;
  p.world = world;
  p.causal = causal;
  p.zigzag = zigzag;
  p.k_batch_stride = k_batch_stride;
  p.scale_log2 = sm_scale * 1.4426950408889634f;
  static bool attr = false;
  if (!attr) {
    TDIST_RT_TRY(cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnSmem::kTotal));
    attr = true;
  }
  dim3 grid((max_q_local + kAtBM - 1) / kAtBM, Hq, B);
  attn_fwd_tc_kernel<<<grid, kAtThreads, AttnSmem::kTotal, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("attn_fwd_tc");
}
