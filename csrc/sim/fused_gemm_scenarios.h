// fused_gemm_scenarios.h — the communication side of the PRODUCTION fused GEMM kernel (csrc/kernels/gemm_sm100.cuh) on the
// simulator. The tensor-core main loop (TMA, mbarrier, tcgen05) cannot run on a CPU; what runs here, per GEMM CTA, is
//   warp 2  the real reducer                 rs_pull_body / ar_reduce_body (called exactly as the kernel calls them)
//   warp 3  the real publisher               the text between the [sim:..._publish] markers of the kernel, pasted in by
//                                            the staging step (its spin on the shared tile counter gets a yield)
//   warp 4  a stand-in for the epilogue      walks the kernel's own tile order (tile_coord), stores each "partial tile" —
//                                            small integers instead of a product — where the epilogue stores it, then
//                                            counts it in shared memory as the epilogue does
// with the kernel's prologue arithmetic (tiles_m, tiles_n, epoch, parity) restated around them. So the flag positions, the
// epochs, the parity halves, the owner-interleaved order and the reduction units are the production ones.
#pragma once

namespace {

float fused_val(int rank, int row, int col) { return float((rank * 5 + row * 3 + col) % 8); }

template <int BLOCK_N>
void fused_cta(const tdist::GemmParams& p) {
  using namespace tdist;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int tiles_m = (p.mode == GEMM_AG || p.mode == GEMM_RS) ? blocks_per_shard(p) * p.ctx.world : (p.M + kBlockM - 1) / kBlockM;
  const int tiles = tiles_m * tiles_n;
  const int num_gemm = gridDim.x - p.num_comm_ctas;
  const int gemm_cta = static_cast<int>(blockIdx.x) - p.num_comm_ctas;
  const uint32_t phase_base = phase_begin(p.ps);
  const uint32_t epoch = phase_base + 2;
  const int parity = (epoch >> 1) & 1;
  typedef uint32_t stored_count_sim_t;
  SIM_SHARED(stored_count_sim_t, stored_count_word);
  uint32_t* stored_count = &stored_count_word;
  if (threadIdx.x == 0) *stored_count = 0;
  __syncthreads();
  if (warp == 4) {
    if (lane == 0) {  // epilogue stand-in
      for (int t = gemm_cta; t < tiles; t += num_gemm) {
        const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(p.c) + size_t(parity) * p.ws_parity_bytes);
        const long long ld = p.mode == GEMM_RS ? p.N : p.ldc;
        const int col0 = tc.n_blk * BLOCK_N;
        for (int r = tc.row0; r < min(tc.row0 + kBlockM, tc.row_end); ++r)
          for (int c = col0; c < min(col0 + BLOCK_N, p.N); ++c) dst[size_t(r) * ld + c] = __float2bfloat16(fused_val(p.ctx.rank, r, c));
        atomicAdd(stored_count, 1u);
        sim_yield();
      }
    }
  } else if (warp == 3) {
    if (lane == 0) {
      if (p.mode == GEMM_RS) {
#include "kernels/gemm_rs_publish.inc"
      } else {
#include "kernels/gemm_ar_publish.inc"
      }
    }
  } else if (warp == 2) {
    if (p.mode == GEMM_RS)
      rs_pull_body<BF16, BLOCK_N>(p, epoch, gemm_cta, num_gemm, tiles_n, false);
    else
      ar_reduce_body<BF16, BLOCK_N>(p, epoch, gemm_cta, num_gemm, tiles_m, tiles_n);
  }
  __syncthreads();
  phase_end(p.ps, phase_base, 2);
}

// GEMM + RS without a multicast mapping (the scatter form): the epilogue stores every partial tile into the OWNER's slot
// buffer over the peer mapping and publishes it one tile later through the kernel's own `publish` lambda (text between the
// [sim:epi_publish] markers: per-(owner, block) tile counter, then the flag on the owner); after its tiles every CTA runs the
// real rs_reduce_body over this rank's segment.
template <int BLOCK_N>
void fused_rs_scatter_cta(const tdist::GemmParams& p) {
  using namespace tdist;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int tiles_m = blocks_per_shard(p) * p.ctx.world;
  const int tiles = tiles_m * tiles_n;
  const int num_gemm = gridDim.x - p.num_comm_ctas;
  const int gemm_cta = static_cast<int>(blockIdx.x) - p.num_comm_ctas;
  const uint32_t phase_base = phase_begin(p.ps);
  const uint32_t epoch = phase_base + 2;
  const int parity = (epoch >> 1) & 1;
  if (warp == 4 && lane == 0) {
#include "kernels/gemm_epi_publish.inc"
    int pend_t = -1, pend_shard = 0, pend_blk = 0, pend_nblk = 0;
    for (int t = gemm_cta; t < tiles; t += num_gemm) {
      const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
      if (pend_t >= 0) publish(pend_t, pend_shard, pend_blk, pend_nblk);  // (the kernel publishes a tile one tile later)
      const size_t seg_elems = size_t(p.rows_per_rank) * p.N;
      __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(symm_at(p.ctx, static_cast<char*>(p.c), tc.shard) + size_t(parity) * p.ws_parity_bytes +
                                                             size_t(p.ctx.rank) * seg_elems * 2);
      const int row0 = tc.row0 - tc.shard * p.rows_per_rank, col0 = tc.n_blk * BLOCK_N;
      for (int r = row0; r < min(row0 + kBlockM, p.rows_per_rank); ++r)
        for (int c = col0; c < min(col0 + BLOCK_N, p.N); ++c)
          dst[size_t(r) * p.N + c] = __float2bfloat16(fused_val(p.ctx.rank, tc.shard * p.rows_per_rank + r, c));
      pend_t = t, pend_shard = tc.shard, pend_blk = tc.blk, pend_nblk = tc.n_blk;
      sim_yield();
    }
    if (pend_t >= 0) publish(pend_t, pend_shard, pend_blk, pend_nblk);
  }
  __syncthreads();
  rs_reduce_body<BF16>(p, epoch);
  phase_end(p.ps, phase_base, 2);
}

// The CTA-pair kernel (gemm_sm100_2cta.cuh, the default of the flagship): two consecutive CTAs share a 256-row tile, CTA
// `cta_rank` of the pair owns its 128-row half; both run a reducer and a publisher of their own.
template <int BLOCK_N>
void fused_pair_cta(const tdist::GemmParams& p) {
  using namespace tdist;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = blockIdx.x & 1u;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int bps128 = blocks_per_shard(p);
  const int ptiles_m = ((p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM)) * p.ctx.world;
  const int tiles = ptiles_m * tiles_n;
  const int num_pairs = (gridDim.x - p.num_comm_ctas) / 2;
  const int gemm_cta = static_cast<int>(blockIdx.x) - p.num_comm_ctas;
  const int pair = gemm_cta >> 1;
  const uint32_t phase_base = phase_begin(p.ps);
  const uint32_t epoch = phase_base + 2;
  const int parity = (epoch >> 1) & 1;
  typedef uint32_t stored_count_sim_t;
  SIM_SHARED(stored_count_sim_t, stored_count_word);
  uint32_t* stored_count = &stored_count_word;
  if (threadIdx.x == 0) *stored_count = 0;
  __syncthreads();
  if (warp == 4) {
    if (lane == 0) {  // epilogue stand-in: my 128-row half of every pair tile, masked by the end of the shard as the stores are
      for (int t = pair; t < tiles; t += num_pairs) {
        const PairCoord tc = pair_coord(p, t, ptiles_m, tiles_n);
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(p.c) + size_t(parity) * p.ws_parity_bytes);
        const int row0 = tc.row0 + int(cta_rank) * kBlockM, col0 = tc.n_blk * BLOCK_N;
        for (int r = row0; r < min(row0 + kBlockM, tc.row_end); ++r)
          for (int c = col0; c < min(col0 + BLOCK_N, p.N); ++c) dst[size_t(r) * p.N + c] = __float2bfloat16(fused_val(p.ctx.rank, r, c));
        atomicAdd(stored_count, 1u);
        sim_yield();
      }
    }
  } else if (warp == 3) {
    if (lane == 0) {
#include "kernels/gemm_rs_publish_pair.inc"
    }
  } else if (warp == 2) {
    const int ncw = static_cast<int>(gridDim.x) - p.num_comm_ctas;
    rs_pull_body<BF16, BLOCK_N>(p, epoch, gemm_cta, ncw, tiles_n, true);
  }
  __syncthreads();
  phase_end(p.ps, phase_base, 2);
}

// AG + GEMM, the in-switch broadcast form (the flagship's first half at 3+ ranks): comm CTAs run the real ag_push_mc_body;
// warp 0 of every GEMM CTA walks the kernel's tile order and runs the producer warp's arrival wait — the text between the
// [sim:ag_wait] markers — in front of a stand-in for the TMA loads of a k block: plain reads of the 128 x 64 A elements the
// loads would fetch (the rows of the tile that belong to its shard), summed per tile for the comparison with the host.
template <int BLOCK_N>
void fused_ag_cta(const tdist::GemmParams& p, unsigned long long* sums) {
  using namespace tdist;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int tiles_m = blocks_per_shard(p) * p.ctx.world;
  const int tiles = tiles_m * tiles_n;
  const int num_k = (p.K + kBlockK - 1) / kBlockK;
  const int num_gemm = gridDim.x - p.num_comm_ctas;
  const int gemm_cta = static_cast<int>(blockIdx.x) - p.num_comm_ctas;
  const uint32_t phase_base = phase_begin(p.ps);
  const uint32_t epoch = phase_base + 2;
  const int parity = (epoch >> 1) & 1;
  if (gemm_cta < 0) {
    if (p.variant == VARIANT_NVLS) {
      ag_push_mc_body(p, epoch, blockIdx.x, p.num_comm_ctas);
    } else {
      SIM_DYN_SHARED(uint8_t, smem);
      ag_push_body(p, smem, epoch, blockIdx.x, p.num_comm_ctas);
    }
  } else if (warp == 0) {
    const int bps = blocks_per_shard(p);
    for (int t = gemm_cta; t < tiles; t += num_gemm) {
      const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
      unsigned long long sum = 0;
#include "kernels/gemm_ag_wait.inc"
        if (lane == 0) {  // stand-in for the loads of k block kb
          const bool local = ta == &p.tma_a_local;
          const uint16_t* src = local ? static_cast<const uint16_t*>(p.a_local)
                                      : reinterpret_cast<const uint16_t*>(static_cast<const char*>(p.a_ws) + size_t(parity) * p.ws_parity_bytes);
          const long long ld = local ? p.a_lda : p.K;
          const int rows = min(kBlockM, tc.row_end - tc.row0);
          for (int r = 0; r < rows; ++r)
            for (int k = kb * kBlockK; k < min((kb + 1) * kBlockK, p.K); ++k) sum += src[size_t(a_row + r) * ld + k];
        }
      }  // k blocks
      __syncwarp();
      if (lane == 0) sums[t] = sum;
    }
  }
  __syncthreads();
  phase_end(p.ps, phase_base, 2);
}

// the same for the CTA-pair kernel (gemm_sm100_2cta.cuh): both CTAs of a pair walk the 256-row pair tiles, each waits for and
// reads its own 128-row half
template <int BLOCK_N>
void fused_ag_pair_cta(const tdist::GemmParams& p, unsigned long long* sums) {
  using namespace tdist;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = blockIdx.x & 1u;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int bps128 = blocks_per_shard(p);
  const int ptiles_m = ((p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM)) * p.ctx.world;
  const int tiles = ptiles_m * tiles_n;
  const int num_k = (p.K + kBlockK - 1) / kBlockK;
  const int num_pairs = (gridDim.x - p.num_comm_ctas) / 2;
  const int gemm_cta = static_cast<int>(blockIdx.x) - p.num_comm_ctas;
  const int pair = gemm_cta >> 1;
  const uint32_t phase_base = phase_begin(p.ps);
  const uint32_t epoch = phase_base + 2;
  const int parity = (epoch >> 1) & 1;
  if (gemm_cta < 0) {
    if (p.variant == VARIANT_NVLS) {
      ag_push_mc_body(p, epoch, blockIdx.x, p.num_comm_ctas);
    } else {
      SIM_DYN_SHARED(uint8_t, smem);
      ag_push_body(p, smem, epoch, blockIdx.x, p.num_comm_ctas);
    }
  } else if (warp == 0) {
    for (int t = pair; t < tiles; t += num_pairs) {
      const PairCoord tc = pair_coord(p, t, ptiles_m, tiles_n);
      unsigned long long sum = 0;
#include "kernels/gemm_ag_wait_pair.inc"
        (void)b_row;
        if (lane == 0) {  // stand-in for the loads of k block kb: my half of the pair tile, up to the end of the shard
          const bool local = ta == &p.tma_a_local;
          const uint16_t* src = local ? static_cast<const uint16_t*>(p.a_local)
                                      : reinterpret_cast<const uint16_t*>(static_cast<const char*>(p.a_ws) + size_t(parity) * p.ws_parity_bytes);
          const long long ld = local ? p.a_lda : p.K;
          const int rows = min(kBlockM, tc.row_end - (tc.row0 + int(cta_rank) * kBlockM));
          for (int r = 0; r < rows; ++r)
            for (int k = kb * kBlockK; k < min((kb + 1) * kBlockK, p.K); ++k) sum += src[size_t(a_row + r) * ld + k];
        }
      }  // k blocks
      __syncwarp();
      if (lane == 0) sums[size_t(t) * 2 + cta_rank] = sum;
    }
  }
  __syncthreads();
  phase_end(p.ps, phase_base, 2);
}

// rows: rows per rank (a multiple of 8 is not required; blocks of 128), `alternate` as below (the smaller calls use half the rows);
// slices == 0 selects the point-to-point form (ag_push_body: bulk-copy ring to every peer, 4 flagged 32-row parts per block)
int run_fused_ag(int world, int max_rows, int N, int K, int push_ctas, int ctas, int calls, int alternate, int slices, int slabs, int fixed_rows, bool pair) {
  using namespace tdist;
  using bf16 = __nv_bfloat16;
  constexpr int BLOCK_N = 64;
  const bool ring = slices == 0;  // the point-to-point form (default with one peer): bulk-copy ring, 4 flagged parts per block
  if (ring) slices = 1;
  if (world < 2 || K % 8 != 0 || slices < 1 || (slices & (slices - 1)) || slices > 32 || push_ctas < 1 || ctas <= push_ctas) {
    fprintf(stderr, "fused ag: needs 2+ ranks (one rank runs the plain GEMM), K %% 8 == 0, a power-of-two slice count <= 32, push CTAs < CTAs\n");
    return 2;
  }
  if (pair && (push_ctas % 2 != 0 || (ctas - push_ctas) % 2 != 0)) {
    fprintf(stderr, "fused ag2: the CTA-pair form needs even numbers of push and GEMM CTAs\n");
    return 2;
  }
  const size_t W = world;
  const size_t parity_bytes = W * max_rows * size_t(K) * 2;
  const int bps_max = (max_rows + kBlockM - 1) / kBlockM;
  const size_t flag_words = W * bps_max * kAgSlots;
  sim::Node node(world, 2 * parity_bytes + flag_words * 4, true);
  auto rows_of = [=](int call) { return alternate && (call & 1) == (alternate & 1) ? max_rows / 2 : max_rows; };
  std::vector<std::vector<bf16>> A(size_t(calls) * world);  // one input per (call, rank): never rewritten
  for (int c = 0; c < calls; ++c)
    for (int r = 0; r < world; ++r) A[size_t(c) * world + r] = random_bf16(size_t(rows_of(c)) * K, 4000 * c + 17 * r + 3, 1.f);
  sim::run_ranks(world, [&](int rank) {
    uint32_t state[2] = {0, 0};
    for (int call = 0; call < calls; ++call) {
      const int rows = rows_of(call);
      GemmParams p;
      memset(&p, 0, sizeof(p));
      p.ctx = node.ctx(rank);
      p.mode = GEMM_AG;
      p.variant = ring ? VARIANT_LEGACY : VARIANT_NVLS;
      p.is_bf16 = 1;
      p.group_m = 8;
      p.M = rows * world;
      p.N = N;
      p.K = K;
      p.rows_per_rank = rows;
      p.a_local = A[size_t(call) * world + rank].data();
      p.a_lda = K;
      p.a_ws = node.base(rank);
      p.ws_parity_bytes = parity_bytes;
      p.flags = reinterpret_cast<uint32_t*>(node.base(rank) + 2 * parity_bytes);
      p.ps = PhaseState{state, state + 1};
      p.num_comm_ctas = push_ctas;
      p.flag_row = fixed_rows ? int(flag_words / W) : 0;  // TDIST_FIXED_FLAG_LAYOUT=1 of csrc/kernels/gemm.cu
      // the host code's split of a block into K slabs (csrc/kernels/gemm.cu)
      const int num_kb = (K + kBlockK - 1) / kBlockK;
      int sl = slabs;
      if (sl > kAgSlots / slices) sl = kAgSlots / slices;
      if (sl > num_kb) sl = num_kb;
      if (sl < 1) sl = 1;
      p.ag_slices = slices;
      p.ag_slab_kblocks = (num_kb + sl - 1) / sl;
      p.ag_kslabs = (num_kb + p.ag_slab_kblocks - 1) / p.ag_slab_kblocks;
      const int tiles_n = (N + BLOCK_N - 1) / BLOCK_N;
      const int tiles_m = pair ? ((rows + 2 * kBlockM - 1) / (2 * kBlockM)) * world : blocks_per_shard(p) * world, tiles = tiles_m * tiles_n;
      std::vector<unsigned long long> sums(size_t(tiles) * (pair ? 2 : 1), ~0ull);
      unsigned long long* sp = sums.data();
      const GemmParams* pp = &p;
      if (pair)
        sim::launch(rank, ctas, 64, [=] { fused_ag_pair_cta<BLOCK_N>(*pp, sp); });
      else
        sim::launch(rank, ctas, 64, [=] { fused_ag_cta<BLOCK_N>(*pp, sp); });
      for (int t = 0; t < tiles; ++t)
        for (int half = 0; half < (pair ? 2 : 1); ++half) {
          int shard, first;
          if (pair) {
            const PairCoord tc = pair_coord(p, t, tiles_m, tiles_n);
            shard = tc.shard;
            first = tc.pblk * 2 * kBlockM + half * kBlockM;
          } else {
            const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
            shard = tc.shard;
            first = tc.blk * kBlockM;
          }
          const uint16_t* src = reinterpret_cast<const uint16_t*>(A[size_t(call) * world + shard].data());
          unsigned long long want = 0;
          for (int r = first; r < std::min(first + kBlockM, rows); ++r)
            for (int k = 0; k < K; ++k) want += src[size_t(r) * K + k];
          const unsigned long long got = sums[pair ? size_t(t) * 2 + half : size_t(t)];
          if (got != want) mismatch(pair ? "fused ag2" : "fused ag", rank, call, t, float(got % 100000), float(want % 100000));
        }
    }
  });
  printf("protocol_sim fused %s: %d ranks x %d calls, %d rows per rank%s, N %d, K %d, %d push + %d GEMM CTAs, %s (%d row slices x %d K slabs per block), flag rows %s: %s\n",
         pair ? "ag2" : "ag", world, calls, max_rows, alternate ? " (alternating with half)" : "", N, K, push_ctas, ctas - push_ctas,
         ring ? "point-to-point ring" : "in-switch broadcast", slices, slabs, fixed_rows ? "fixed" : "per call", g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

// mode: "rs" GEMM + ReduceScatter (NVLS pull protocol; "rsl": the scatter form used without a multicast mapping; "rs2": the CTA-pair kernel's form, `ctas` must be even; `mc` selects in-switch reduction or peer loads), "ar1" / "ar2" GEMM +
// AllReduce one-shot / two-shot. Calls alternate between M and M / 2 on the same buffers when `alternate` is 1 (odd calls
// smaller) or 2 (even calls smaller). `fixed_rows`: one flag row per source rank of constant length (GemmParams::flag_row).
int run_fused(const std::string& mode, int world, int max_M, int N, int ctas, int calls, int mc, int alternate, int fixed_rows) {
  if (mode == "rs2" && (ctas < 2 || ctas % 2 != 0)) {
    fprintf(stderr, "fused rs2: the CTA-pair form needs an even number of CTAs\n");
    return 2;
  }
  using namespace tdist;
  using bf16 = __nv_bfloat16;
  constexpr int BLOCK_N = 64;
  const bool pair = mode == "rs2", scatter = mode == "rsl", rs = mode == "rs" || pair || scatter, one_shot = mode == "ar1";
  const size_t W = world;
  const size_t parity_bytes = size_t(max_M) * N * 2;
  const size_t out_bytes = size_t(max_M) * N * 2;  // AR two-shot: the output is symmetric (peers store into it)
  const size_t flag_words = W * size_t((max_M + 127) / 128 + W) * ((N + BLOCK_N - 1) / BLOCK_N) + 64;
  const size_t off_out = 2 * parity_bytes, off_flags = off_out + out_bytes, off_flags2 = off_flags + flag_words * 4;
  sim::Node node(world, off_flags2 + (TDIST_MAX_RANKS + 64) * 4, mc != 0);
  auto M_of = [=](int call) { return alternate && (call & 1) == (alternate & 1) ? max_M / 2 : max_M; };
  sim::run_ranks(world, [&](int rank) {
    uint32_t state[2] = {0, 0};
    std::vector<uint32_t> counter_words(W * size_t((max_M / world + 127) / 128) + 4, 0);  // scatter form: tiles per (owner, block)
    uint32_t* counters = counter_words.data();
    std::vector<bf16> out_local(size_t(max_M) * N);
    for (int call = 0; call < calls; ++call) {
      const int M = M_of(call);
      GemmParams p;
      memset(&p, 0, sizeof(p));
      p.ctx = node.ctx(rank);
      p.c = node.base(rank);
      p.ws_parity_bytes = parity_bytes;
      p.M = M;
      p.N = N;
      p.K = 64;
      p.ldc = N;
      p.mode = rs ? GEMM_RS : GEMM_AR;
      p.variant = scatter ? VARIANT_LEGACY : VARIANT_NVLS;
      p.is_bf16 = 1;
      p.group_m = 8;
      p.rows_per_rank = rs ? M / world : M;
      p.flags = reinterpret_cast<uint32_t*>(node.base(rank) + off_flags);
      p.flags2 = reinterpret_cast<uint32_t*>(node.base(rank) + off_flags2);
      p.counters = counters;
      p.ps = PhaseState{state, state + 1};
      p.ar_one_shot = one_shot;
      p.use_mc = mc != 0 && node.mc_window != nullptr;  // the host code's rule (a single rank has no multicast mapping)
      p.rs_unit_rows = 16;
      p.flag_row = fixed_rows ? int(flag_words / W) : 0;  // TDIST_FIXED_FLAG_LAYOUT=1 of csrc/kernels/gemm.cu
      // RS and one-shot AR write a local result; two-shot AR writes the symmetric output of every rank
      p.out = (rs || one_shot) ? static_cast<void*>(out_local.data()) : static_cast<void*>(node.base(rank) + off_out);
      const GemmParams* pp = &p;
      if (pair)
        sim::launch(rank, ctas, 256, [=] { fused_pair_cta<BLOCK_N>(*pp); });
      else if (scatter)
        sim::launch(rank, ctas, 256, [=] { fused_rs_scatter_cta<BLOCK_N>(*pp); });
      else
        sim::launch(rank, ctas, 256, [=] { fused_cta<BLOCK_N>(*pp); });
      const bf16* got = static_cast<const bf16*>(p.out);
      const int rows = rs ? M / world : M, row_base = rs ? rank * (M / world) : 0;
      for (int r = 0; r < rows; ++r)
        for (int c = 0; c < N; ++c) {
          float want = 0.f;
          for (size_t q = 0; q < W; ++q) want += fused_val(int(q), row_base + r, c);
          const float g = __bfloat162float(got[size_t(r) * N + c]);
          if (g != want) mismatch(("fused " + mode).c_str(), rank, call, long(r) * N + c, g, want);
        }
      if (!rs && !one_shot) {
        // the caller of the two-shot op reads `out` and the next call overwrites it on every rank: keep the ranks together,
        // as the layer above does by consuming the result before it launches again
        uint32_t* bflags = p.flags2 + 32;
        const SymmCtx cx = p.ctx;
        const PhaseState ps = p.ps;
        sim::launch(rank, 1, 32, [=] { barrier_all_kernel(cx, bflags, ps); });
      }
    }
  });
  printf("protocol_sim fused %s: %d ranks x %d calls, M %d%s N %d, %d GEMM CTAs, %s, flag rows %s: %s\n", mode.c_str(), world, calls, max_M,
         alternate ? " (alternating with half)" : "", N, ctas, mc ? "in-switch reduction" : "peer loads", fixed_rows ? "fixed" : "per call",
         g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

}  // namespace
