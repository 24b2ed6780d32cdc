// gemm_sm100_2cta.cuh — CTA-pair (cta_group::2) variant of the tcgen05 GEMM engine.
//
// Two CTAs of a cluster (same TPC) compute one 256 x BLOCK_N tile: each CTA TMA-loads its own 128 rows of A and HALF of
// the B tile, the leader CTA issues tcgen05.mma.cta_group::2 (UMMA M = 256) which reads A/B from both CTAs' shared
// memory, and each CTA's TMEM receives the accumulator of its own 128 rows. Per CTA this halves the B traffic into
// shared memory (TMA writes + MMA operand reads), which is what limits the 1-CTA kernel at 128 x N tiles — narrow
// tiles (needed against wave quantisation when N/W is small) stay efficient. Modes: PLAIN, AG, RS (same protocols
// and the same comm / reduce bodies as gemm_sm100.cuh).
//
// Barrier wiring (per smem stage unless noted):
//   full  (leader CTA only, count 2): leader producer arrive.expect_tx(bytes of BOTH CTAs) + peer producer remote arrive;
//                                     both CTAs' TMA loads complete_tx on the leader's barrier (cta_group::2 TMA form)
//   empty (each CTA, count 1)       : leader's tcgen05.commit multicast to both CTAs
//   tmem_full  (each CTA, count 1)  : leader's tcgen05.commit multicast            (per accumulator stage)
//   tmem_empty (leader only, count 8): 4 epilogue warps of each CTA (peer arrives remotely)
#pragma once
#include "gemm_sm100.cuh"

namespace tdist {

constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared::cluster address -> leader CTA

template <int BLOCK_N, int STAGES>
struct GemmSmem2 {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = (BLOCK_N / 2) * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiBytes = 4 * 32 * 128;
  static constexpr int kBarOffset = STAGES * kStageBytes + kEpiBytes;
  static constexpr int kGemmTotal = kBarOffset + (2 * STAGES + 4) * 8 + 16;
  static constexpr int kPushTotal = 6 * 32 * 1024 + 64;
  static constexpr int kTotal = kGemmTotal > kPushTotal ? kGemmTotal : kPushTotal;
};

// tile order in units of 256-row pair blocks (same policy as tile_coord in gemm_sm100.cuh)
struct PairCoord {
  int shard, pblk, row0, row_end, n_blk;
};
__device__ __forceinline__ PairCoord pair_coord(const GemmParams& p, int t, int ptiles_m, int tiles_n) {
  PairCoord tc;
  const int gm = max(1, p.group_m / 2);
  if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
    // owner-interleaved walk, identical on every rank (see tile_coord)
    const int W = p.ctx.world;
    const int bps2 = (p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM);
    tc.shard = t % W;
    const int i = t / W;
    tc.pblk = i % bps2;
    tc.n_blk = i / bps2;
    tc.row0 = tc.shard * p.rows_per_rank + tc.pblk * 2 * kBlockM;
    tc.row_end = (tc.shard + 1) * p.rows_per_rank;
  } else if (p.mode == GEMM_AG && p.variant == VARIANT_NVLS &&
             t >= ((p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM)) * tiles_n) {
    // remote shards block-major (see tile_coord)
    const int W = p.ctx.world;
    const int bps2 = (p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM);
    const int r = t - bps2 * tiles_n;
    const int unit = r / tiles_n;
    tc.n_blk = r - unit * tiles_n;
    tc.pblk = unit / (W - 1);
    tc.shard = (p.ctx.rank + 1 + unit % (W - 1)) % W;
    tc.row0 = tc.shard * p.rows_per_rank + tc.pblk * 2 * kBlockM;
    tc.row_end = (tc.shard + 1) * p.rows_per_rank;
  } else if (p.mode == GEMM_AG || p.mode == GEMM_RS) {
    const int W = p.ctx.world;
    const int bps2 = (p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM);
    const int per_shard = bps2 * tiles_n;
    const int step = t / per_shard;
    const int r = t - step * per_shard;
    tc.shard = (p.ctx.rank + step + (p.mode == GEMM_RS ? 1 : 0)) % W;
    const int per_group = gm * tiles_n;
    const int g = r / per_group;
    const int first_m = g * gm;
    const int rows = min(gm, bps2 - first_m);
    const int rr = r - g * per_group;
    tc.pblk = first_m + rr % rows;
    tc.n_blk = rr / rows;
    tc.row0 = tc.shard * p.rows_per_rank + tc.pblk * 2 * kBlockM;
    tc.row_end = (tc.shard + 1) * p.rows_per_rank;
  } else {
    const int per_group = gm * tiles_n;
    const int g = t / per_group;
    const int first_m = g * gm;
    const int rows = min(gm, ptiles_m - first_m);
    const int r = t - g * per_group;
    tc.shard = 0;
    tc.pblk = first_m + r % rows;
    tc.n_blk = r / rows;
    tc.row0 = tc.pblk * 2 * kBlockM;
    tc.row_end = p.M;
  }
  return tc;
}

template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_sm100_2cta_kernel(const __grid_constant__ GemmParams p) {
  using L = GemmSmem2<BLOCK_N, STAGES>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = ptx::cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int bps128 = blocks_per_shard(p);
  const int ptiles_m = (p.mode == GEMM_AG || p.mode == GEMM_RS)
                           ? ((p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM)) * p.ctx.world
                           : (p.M + 2 * kBlockM - 1) / (2 * kBlockM);
  const int tiles = ptiles_m * tiles_n;
  const int num_k = (p.K + kBlockK - 1) / kBlockK;
  const int num_pairs = (gridDim.x - p.num_comm_ctas) / 2;
  const int gemm_cta = static_cast<int>(blockIdx.x) - p.num_comm_ctas;
  const int pair = gemm_cta >> 1;

  uint32_t epoch = 0, phase_base = 0;
  if (p.mode != GEMM_PLAIN) {
    phase_base = phase_begin(p.ps);
    epoch = phase_base + 2;
  }
  const int parity = (epoch >> 1) & 1;

  if (gemm_cta < 0) {
    if (p.mode == GEMM_AG) {
      if (p.variant == VARIANT_NVLS)
        ag_push_mc_body(p, epoch, blockIdx.x, p.num_comm_ctas);
      else
        ag_push_body(p, smem, epoch, blockIdx.x, p.num_comm_ctas);
    }
  } else {
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * L::kABytes;
    uint8_t* smem_epi = smem + STAGES * L::kStageBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full = empty_bar + STAGES;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    uint32_t* stored_count = tmem_ptr_smem + 1;  // RS (NVLS): tiles of this CTA whose partial result is in memory
    constexpr int kTmemCols = tmem_cols<BLOCK_N>();
    constexpr int kAccStride = (BLOCK_N + 31) / 32 * 32;

    if (warp == 0 && lane == 0) {
      ptx::prefetch_tensormap(&p.tma_a[0]);
      ptx::prefetch_tensormap(&p.tma_b);
    }
    if (warp == 1 && lane == 0) {
      for (int i = 0; i < STAGES; ++i) {
        ptx::mbar_init(full_bar + i, 2);
        ptx::mbar_init(empty_bar + i, 1);
      }
      for (int i = 0; i < 2; ++i) {
        ptx::mbar_init(tmem_full + i, 1);
        ptx::mbar_init(tmem_empty + i, 8);
      }
      *stored_count = 0;
      ptx::fence_barrier_init();
    }
    if (warp == 2) ptx::tmem_alloc_cg2<kTmemCols>(tmem_ptr_smem);
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();  // both CTAs' barriers are initialised before anybody arrives remotely
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
      // ======================================================== TMA producer (both CTAs)
      // whole warp walks the tiles (lanes poll the arrival flags of a block in parallel), lane 0 issues the copies
      {
        int stage = 0;
        uint32_t ph = 0;
        for (int t = pair; t < tiles; t += num_pairs) {
          const PairCoord tc = pair_coord(p, t, ptiles_m, tiles_n);
          // [sim:ag_wait_pair begin] (csrc/sim runs the text between these markers against ag_push_mc_body on the CPU)
          const CUtensorMap* ta = &p.tma_a[0];
          int a_row = tc.row0 + int(cta_rank) * kBlockM;  // my 128 rows of the 256-row tile
          const int blk128 = tc.pblk * 2 + int(cta_rank);
          const uint32_t* slab_flags = nullptr;
          if (p.mode == GEMM_AG) {
            if (tc.shard == p.ctx.rank) {
              ta = &p.tma_a_local;
              a_row -= tc.shard * p.rows_per_rank;
            } else {
              ta = &p.tma_a[parity];
              if (blk128 < bps128) {
                if (p.variant != VARIANT_NVLS) {
                  if (lane < kAgParts)
                    wait_ge_sys_poll_relaxed(p.flags + size_t(tc.shard) * (p.flag_row > 0 ? p.flag_row : bps128 * kAgParts) + blk128 * kAgParts + lane, epoch);
                  __syncwarp();
                  ptx::fence_proxy_async();
                } else {
                  slab_flags = p.flags + size_t(tc.shard) * (p.flag_row > 0 ? p.flag_row : bps128 * kAgSlots) + blk128 * kAgSlots;
                }
              }
            }
          }
          const int b_row = tc.n_blk * BLOCK_N + int(cta_rank) * (BLOCK_N / 2);
          for (int kb = 0; kb < num_k; ++kb) {
            if (slab_flags != nullptr && kb % p.ag_slab_kblocks == 0) {
              // NVLS push: the K loop follows the arrival of my 128 rows' K slabs
              if (lane < p.ag_slices) wait_ge_sys_poll_relaxed(slab_flags + (kb / p.ag_slab_kblocks) * p.ag_slices + lane, epoch);
              __syncwarp();
              ptx::fence_proxy_async();
            }
            // [sim:ag_wait_pair end]
            if (lane == 0) {
              ptx::mbar_wait(empty_bar + stage, ph ^ 1);
              const uint32_t leader_full = ptx::smem_u32(full_bar + stage) & kPeerBitMask;
              if (leader)
                ptx::mbar_arrive_expect_tx(full_bar + stage, 2 * L::kStageBytes);
              else
                ptx::mbar_arrive_cluster(full_bar + stage, 0);
              ptx::tma_load_2d_cg2(smem_a + stage * L::kABytes, ta, leader_full, kb * kBlockK, a_row);
              ptx::tma_load_2d_cg2(smem_b + stage * L::kBBytes, &p.tma_b, leader_full, kb * kBlockK, b_row);
            }
            if (++stage == STAGES) {
              stage = 0;
              ph ^= 1;
            }
          }
          __syncwarp();
        }
      }
    } else if (warp == 1) {
      // ======================================================== MMA issuer (leader CTA only)
      if (leader && lane == 0) {
        const uint32_t idesc = ptx::make_idesc_f16(2 * kBlockM, BLOCK_N, p.is_bf16 != 0);
        int stage = 0;
        uint32_t ph = 0;
        int acc = 0;
        uint32_t acc_ph = 0;
        for (int t = pair; t < tiles; t += num_pairs) {
          ptx::mbar_wait(tmem_empty + acc, acc_ph ^ 1);
          ptx::tcgen05_fence_after();
          const uint32_t d_tmem = tmem_base + acc * kAccStride;
          for (int kb = 0; kb < num_k; ++kb) {
            ptx::mbar_wait(full_bar + stage, ph);
            ptx::tcgen05_fence_after();
            const uint64_t da = ptx::make_smem_desc_sw128(smem_a + stage * L::kABytes);
            const uint64_t db = ptx::make_smem_desc_sw128(smem_b + stage * L::kBBytes);
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k)
              ptx::umma_f16_ss_cg2(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            ptx::umma_commit_cg2(empty_bar + stage, 0b11);
            if (++stage == STAGES) {
              stage = 0;
              ph ^= 1;
            }
          }
          ptx::umma_commit_cg2(tmem_full + acc, 0b11);
          if (++acc == 2) {
            acc = 0;
            acc_ph ^= 1;
          }
        }
      }
    } else if (warp >= 4) {
      // ======================================================== epilogue (both CTAs, own 128 rows)
      const int q = warp - 4;
      uint8_t* stg = smem_epi + q * (32 * 128);
      int acc = 0;
      uint32_t acc_ph = 0;
      const int W = p.ctx.world;
      int pend_shard = 0, pend_blk = -1, pend_nblk = 0;
      auto publish = [&](int shard, int blk, int n_blk) {
        if (!(p.debug & 2)) ptx::fence_acq_rel_sys();
        if (p.variant == VARIANT_NVLS) {
          // my 128 rows of the partial tile are in my own memory: tell the owner they may be summed
          notify(p.ctx, p.flags + (size_t(p.ctx.rank) * bps128 + blk) * tiles_n + n_blk, shard, epoch);
          return;
        }
        const int cidx = shard * bps128 + blk;
        const uint32_t done = ptx::atom_add_acq_rel_gpu(p.counters + cidx, 1) + 1;
        if (done == static_cast<uint32_t>(tiles_n)) {
          p.counters[cidx] = 0;
          ptx::fence_acq_rel_sys();
          notify(p.ctx, p.flags + p.ctx.rank * bps128 + blk, shard, epoch);
        }
      };
      for (int t = pair; t < tiles; t += num_pairs) {
        const PairCoord tc = pair_coord(p, t, ptiles_m, tiles_n);
        ptx::mbar_wait(tmem_full + acc, acc_ph);
        ptx::tcgen05_fence_after();
        if (pend_blk >= 0 && warp == 4 && lane == 0) publish(pend_shard, pend_blk, pend_nblk);
        pend_blk = -1;
        char* dst_base = static_cast<char*>(p.c);
        long long ld = p.ldc;
        int row0 = tc.row0 + int(cta_rank) * kBlockM + q * 32;
        int row_limit = tc.row_end;
        if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
          dst_base += size_t(parity) * p.ws_parity_bytes;
          ld = p.N;
        } else if (p.mode == GEMM_RS) {
          const size_t seg_elems = size_t(p.rows_per_rank) * p.N;
          dst_base = symm_at(p.ctx, static_cast<char*>(p.c), tc.shard) + size_t(parity) * p.ws_parity_bytes +
                     size_t(p.ctx.rank) * seg_elems * 2;
          row0 -= tc.shard * p.rows_per_rank;
          row_limit = p.rows_per_rank;
          ld = p.N;
        }
        const int col0 = tc.n_blk * BLOCK_N;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N; c += 64) {
          uint32_t v0[32], v1[32];
          const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + acc * kAccStride + c;
          ptx::tmem_ld_32x32b_x32(taddr, v0);
          if (c + 32 < BLOCK_N) ptx::tmem_ld_32x32b_x32(taddr + 32, v1);
          ptx::tmem_ld_wait();
          if (c + 64 >= BLOCK_N) {
            ptx::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (leader)
                ptx::mbar_arrive(tmem_empty + acc);
              else
                ptx::mbar_arrive_cluster(tmem_empty + acc, 0);
            }
          }
          uint32_t pk[32];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float a0 = __uint_as_float(v0[2 * i]), a1 = __uint_as_float(v0[2 * i + 1]);
            const float b0 = __uint_as_float(v1[2 * i]), b1 = __uint_as_float(v1[2 * i + 1]);
            pk[i] = p.is_bf16 ? ptx::pack_bf16x2(a0, a1) : ptx::pack_f16x2(a0, a1);
            pk[16 + i] = p.is_bf16 ? ptx::pack_bf16x2(b0, b1) : ptx::pack_f16x2(b0, b1);
          }
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
            if (grow < row_limit && gcol < p.N && c + cj * 8 < BLOCK_N) ptx::st_v4(dst_base + (size_t(grow) * ld + gcol) * 2, w);
          }
        }
        if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
          // the partial tile is in local memory: count it, warp 3 publishes it (the system fence of a publish also
          // waits for the reducer warp's outstanding NVLS loads; see gemm_sm100.cuh)
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (warp == 4 && lane == 0) {
            __threadfence_block();
            atomicAdd(stored_count, 1u);
          }
        } else if (p.mode == GEMM_RS) {
          asm volatile("bar.sync 1, 128;" ::: "memory");
          const int blk128 = tc.pblk * 2 + int(cta_rank);
          if (blk128 < bps128) {
            pend_shard = tc.shard;
            pend_blk = blk128;
            pend_nblk = tc.n_blk;
          }
        }
        if (++acc == 2) {
          acc = 0;
          acc_ph ^= 1;
        }
      }
      if (pend_blk >= 0 && warp == 4 && lane == 0) publish(pend_shard, pend_blk, pend_nblk);
    } else if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
      if (warp == 3) {
        // ====================================================== warp 3: publishes this CTA's finished half tiles
        if (lane == 0) {
          // [sim:rs_publish_pair begin] (run on the CPU against rs_pull_body by csrc/sim, see gemm_sm100.cuh)
          uint32_t k = 0;
          for (int t = pair; t < tiles; t += num_pairs, ++k) {
            const PairCoord tc = pair_coord(p, t, ptiles_m, tiles_n);
            while (*reinterpret_cast<volatile uint32_t*>(stored_count) <= k) {
            }
            __threadfence_block();
            const int blk128 = tc.pblk * 2 + int(cta_rank);
            if (blk128 >= bps128) continue;
            uint32_t* flag = symm_at(
                p.ctx, p.flags + size_t(p.ctx.rank) * (p.flag_row > 0 ? p.flag_row : bps128 * tiles_n) + blk128 * tiles_n + tc.n_blk, tc.shard);
            if (!(p.debug & 8)) {
              // The partial tile is in THIS GPU's memory and the owner's switch reads are served from this GPU's L2,
              // so a GPU-scope fence orders the tile before the flag; the flag itself is a plain system-scope store.
              // (A system-scope fence here costs ~35 us of the 8-rank op: it waits for the NVLS loads of warp 2.)
              ptx::fence_acq_rel_gpu();
              ptx::st_relaxed_sys(flag, epoch);
            } else {
              if (!(p.debug & 2)) ptx::fence_acq_rel_sys();
              ptx::st_release_sys(flag, epoch);
            }
          }
          // [sim:rs_publish_pair end]
        }
      } else {
        // ====================================================== warp 2: NVLS reducer of my own segment
        const int ncw = static_cast<int>(gridDim.x) - p.num_comm_ctas;
        if (p.is_bf16)
          rs_pull_body<BF16, BLOCK_N>(p, epoch, gemm_cta, ncw, tiles_n, true);
        else
          rs_pull_body<F16, BLOCK_N>(p, epoch, gemm_cta, ncw, tiles_n, true);
      }
    }
    // teardown: the pair stays together until both CTAs are done with each other's shared memory / barriers
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();
    if (warp == 2) ptx::tmem_dealloc_cg2<kTmemCols>(tmem_base);
  }

  if (p.mode == GEMM_RS && p.variant == VARIANT_LEGACY) {
    __syncthreads();
    if (p.is_bf16)
      rs_reduce_body<BF16>(p, epoch);
    else
      rs_reduce_body<F16>(p, epoch);
  }
  if (p.mode != GEMM_PLAIN) phase_end(p.ps, phase_base, 2);
}

}  // namespace tdist
