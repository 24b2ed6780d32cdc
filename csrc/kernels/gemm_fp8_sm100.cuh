// gemm_fp8_sm100.cuh — block-scaled fp8 (e4m3) GEMM on tcgen05.mma.kind::f8f6f4, plain and grouped (MoE).
//
//   C[M, N] (bf16) = sum_kb  sa[m, kb] * sb[n / 128, kb] * ( A8[m, kb*128 .. +128] . B8[n, kb*128 .. +128] )
//
// A8 / B8 are e4m3, K-major; sa holds one fp32 scale per row and 128-wide K block ("1 x 128", what the EP dispatch of
// csrc/kernels/ep_a2a.cu delivers with its fp8 payload), sb one per 128 x 128 weight block. The reference has no fp8
// GEMM at all (its grouped GEMMs are bf16/fp16, moe_reduce_rs.py:69; fp8 exists only as the all-to-all payload and the
// tutorial quantiser, tutorials/04-deepseek-infer-all2all.py:334-427): this is the consumer those fp8 rows were missing.
//
// The scales change every 128 k-elements, so partial sums cannot stay in one tensor-core accumulator. Two-level
// accumulation: for every k block the MMA warp issues 4 x (M128 N128 K32) UMMAs into one of two TMEM accumulator stages
// (fresh, no accumulate across k blocks); the four epilogue warps read that partial (tcgen05.ld), multiply it by the
// block's scale and add it to fp32 registers (each thread owns one row x 128 columns), while the tensor core already
// fills the other stage. Warp roles and the TMA / mbarrier pipeline are those of gemm_sm100.cuh.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace tdist {

constexpr int kF8BlockM = 128;
constexpr int kF8BlockN = 128;
constexpr int kF8BlockK = 128;  // one 128-byte swizzle row of e4m3 = one scale block
constexpr int kF8UmmaK = 32;
constexpr int kF8Stages = 6;
constexpr int kF8Threads = 256;

struct Fp8Params {
  CUtensorMap tma_a, tma_b;
  void* c;
  long long ldc;
  int M, N, K;
  const float* sa;   // [M][sa_ld]
  long long sa_ld;   // floats per row (>= K / 128)
  const float* sb;   // [num_experts][ceil(N/128)][ceil(K/128)]
  const int* block_expert;  // grouped: expert of each 128-row block (negative: padding block), nullptr: plain GEMM
  int is_bf16;       // output type: bf16 or fp16
};

struct Fp8Smem {
  static constexpr int kABytes = kF8BlockM * kF8BlockK;
  static constexpr int kBBytes = kF8BlockN * kF8BlockK;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiBytes = 4 * 32 * 128;
  static constexpr int kBarOffset = kF8Stages * kStageBytes + kEpiBytes;
  static constexpr int kTotal = kBarOffset + (2 * kF8Stages + 4) * 8 + 16;
};

__global__ void __launch_bounds__(kF8Threads, 1) gemm_fp8_sm100_kernel(const __grid_constant__ Fp8Params p) {
  using L = Fp8Smem;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m = (p.M + kF8BlockM - 1) / kF8BlockM;
  const int tiles_n = (p.N + kF8BlockN - 1) / kF8BlockN;
  const int tiles = tiles_m * tiles_n;
  const int num_k = (p.K + kF8BlockK - 1) / kF8BlockK;

  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kF8Stages * L::kABytes;
  uint8_t* smem_epi = smem + kF8Stages * L::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty_bar = full_bar + kF8Stages;
  uint64_t* tmem_full = empty_bar + kF8Stages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  constexpr int kTmemCols = 2 * kF8BlockN;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&p.tma_a);
    ptx::prefetch_tensormap(&p.tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kF8Stages; ++i) {
      ptx::mbar_init(full_bar + i, 1);
      ptx::mbar_init(empty_bar + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(tmem_full + i, 1);
      ptx::mbar_init(tmem_empty + i, 4);  // one arrival per epilogue warp
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc<kTmemCols>(tmem_ptr_smem);
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ======================================================== TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t ph = 0;
      for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int m_blk = t / tiles_n, n_blk = t - m_blk * tiles_n;
        int b_row = n_blk * kF8BlockN;
        if (p.block_expert != nullptr) {
          const int e = p.block_expert[m_blk];
          if (e < 0) continue;
          b_row += e * p.N;
        }
        for (int kb = 0; kb < num_k; ++kb) {
          ptx::mbar_wait(empty_bar + stage, ph ^ 1);
          ptx::mbar_arrive_expect_tx(full_bar + stage, L::kStageBytes);
          ptx::tma_load_2d(smem_a + stage * L::kABytes, &p.tma_a, full_bar + stage, kb * kF8BlockK, m_blk * kF8BlockM);
          ptx::tma_load_2d(smem_b + stage * L::kBBytes, &p.tma_b, full_bar + stage, kb * kF8BlockK, b_row);
          if (++stage == kF8Stages) {
            stage = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================== MMA issuer: one fresh partial accumulator per k block
    if (lane == 0) {
      const uint32_t idesc = ptx::make_idesc_f8(kF8BlockM, kF8BlockN, false, false);
      int stage = 0;
      uint32_t ph = 0;
      int acc = 0;
      uint32_t acc_ph = 0;
      for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
        if (p.block_expert != nullptr && p.block_expert[t / tiles_n] < 0) continue;
        for (int kb = 0; kb < num_k; ++kb) {
          ptx::mbar_wait(tmem_empty + acc, acc_ph ^ 1);
          ptx::mbar_wait(full_bar + stage, ph);
          ptx::tcgen05_fence_after();
          const uint32_t d_tmem = tmem_base + acc * kF8BlockN;
          const uint64_t da = ptx::make_smem_desc_sw128(smem_a + stage * L::kABytes);
          const uint64_t db = ptx::make_smem_desc_sw128(smem_b + stage * L::kBBytes);
#pragma unroll
          for (int k = 0; k < kF8BlockK / kF8UmmaK; ++k)  // +32 bytes per K step inside the swizzle row (encoded >> 4)
            ptx::umma_f8f6f4_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, k != 0);
          ptx::umma_commit(empty_bar + stage);  // the smem stage is free when these MMAs retire ...
          ptx::umma_commit(tmem_full + acc);    // ... and the partial accumulator is complete
          if (++stage == kF8Stages) {
            stage = 0;
            ph ^= 1;
          }
          if (++acc == 2) {
            acc = 0;
            acc_ph ^= 1;
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ======================================================== promotion + epilogue: thread = one row x 128 columns
    const int q = warp - 4;
    uint8_t* stg = smem_epi + q * (32 * 128);
    const int kblocks = num_k;
    const int nblk_total = tiles_n;
    int acc = 0;
    uint32_t acc_ph = 0;
    for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
      const int m_blk = t / tiles_n, n_blk = t - m_blk * tiles_n;
      int e = 0;
      if (p.block_expert != nullptr) {
        e = p.block_expert[m_blk];
        if (e < 0) continue;
      }
      const int row = m_blk * kF8BlockM + q * 32 + lane;
      const float* sa_row = p.sa + size_t(row < p.M ? row : 0) * p.sa_ld;
      const float* sb_blk = p.sb + (size_t(e) * nblk_total + n_blk) * kblocks;
      float sum[kF8BlockN];
#pragma unroll
      for (int i = 0; i < kF8BlockN; ++i) sum[i] = 0.f;
      float s_next = row < p.M ? sa_row[0] * sb_blk[0] : 0.f;
      for (int kb = 0; kb < num_k; ++kb) {
        const float s = s_next;
        if (kb + 1 < num_k) s_next = row < p.M ? sa_row[kb + 1] * sb_blk[kb + 1] : 0.f;  // fetched under the TMEM wait
        ptx::mbar_wait(tmem_full + acc, acc_ph);
        ptx::tcgen05_fence_after();
        const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + acc * kF8BlockN;
        uint32_t v0[32], v1[32];
        ptx::tmem_ld_32x32b_x32(taddr, v0);
#pragma unroll
        for (int c = 0; c < kF8BlockN; c += 64) {
          ptx::tmem_ld_32x32b_x32(taddr + c + 32, v1);
          ptx::tmem_ld_wait();  // v0 and v1 of this half have landed
          if (c + 64 >= kF8BlockN) {
            // last TMEM read of this partial: hand the accumulator stage back to the MMA warp
            ptx::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(tmem_empty + acc);
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) sum[c + i] = fmaf(__uint_as_float(v0[i]), s, sum[c + i]);
          if (c + 64 < kF8BlockN) ptx::tmem_ld_32x32b_x32(taddr + c + 64, v0);
#pragma unroll
          for (int i = 0; i < 32; ++i) sum[c + 32 + i] = fmaf(__uint_as_float(v1[i]), s, sum[c + 32 + i]);
        }
        if (++acc == 2) {
          acc = 0;
          acc_ph ^= 1;
        }
      }
      // ---- store the tile: pack to 16 bit, transpose through shared memory, 128-byte coalesced row stores
      const int row0 = m_blk * kF8BlockM + q * 32;
      const int col0 = n_blk * kF8BlockN;
      char* dst_base = static_cast<char*>(p.c);
#pragma unroll
      for (int c = 0; c < kF8BlockN; c += 64) {
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 32; ++i)
          pk[i] = p.is_bf16 ? ptx::pack_bf16x2(sum[c + 2 * i], sum[c + 2 * i + 1]) : ptx::pack_f16x2(sum[c + 2 * i], sum[c + 2 * i + 1]);
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          int4 w = make_int4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          *reinterpret_cast<int4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = w;
        }
        __syncwarp();
        const int cj = lane & 7;
        const int gcol = col0 + c + cj * 8;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + (lane >> 3);
          const int4 w = *reinterpret_cast<const int4*>(stg + r * 128 + ((cj ^ (r & 7)) << 4));
          const int grow = row0 + r;
          if (grow < p.M && gcol < p.N) ptx::st_v4(dst_base + (size_t(grow) * p.ldc + gcol) * 2, w);
        }
      }
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<kTmemCols>(tmem_base);
}

}  // namespace tdist
