// gemm_sm100.cuh — the shared tcgen05/TMEM/TMA GEMM engine.
//
// One persistent, warp-specialised kernel computes C[M,N] = A[M,K] * B[N,K]^T
// (both operands K-major, bf16/fp16 in, fp32 accumulate in TMEM) and is
// specialised at run time into the fused compute+communication ops:
//
//   PLAIN  plain GEMM (gemm_persistent)
//   AG     AllGather+GEMM: extra "comm" CTAs push this rank's A shard to every
//          peer with bulk async copies over NVLink while the GEMM CTAs consume
//          shards in arrival order, gated per 128-row block by flags
//          (reference op: kernels/nvidia/allgather_gemm.py:158-264, 534-653 —
//          there the gather runs on the copy engine of a second stream).
//   RS     GEMM+ReduceScatter: the epilogue stores each partial tile straight
//          into the owner rank's scatter slot over NVLink, a per-block counter
//          publishes completion, and the same kernel finishes with the
//          flag-gated fp32 reduction of the W slots
//          (reference: gemm_reduce_scatter.py:121-250 FUSE_SCATTER + K15).
//   AR     GEMM+AllReduce: tiles go to a symmetric buffer with a per-tile flag;
//          comm CTAs reduce finished tiles across ranks with NVLS
//          multimem.ld_reduce (or P2P loads) while later tiles are computed
//          (reference: gemm_allreduce.py:123-315).
//
// Warp roles per GEMM CTA (256 threads): warp 0 = TMA producer, warp 1 = MMA
// issuer (single elected thread, tcgen05.mma cta_group::1, M=128), warp 2 =
// TMEM allocator, warps 4-7 = epilogue (tcgen05.ld -> bf16 -> smem transpose
// -> coalesced 128-byte row stores, local or peer). Accumulators are double
// buffered in TMEM so the epilogue of tile i overlaps the mainloop of tile i+1.
// Comm CTAs take the LOWEST block indices so they are always scheduled.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace tdist {

enum GemmMode : int { GEMM_PLAIN = 0, GEMM_AG = 1, GEMM_RS = 2, GEMM_AR = 3 };

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // 64 x 16-bit = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 256;
constexpr int kAgParts = 4;  // AG (unicast ring): each 128-row block is pushed as 4 independently flagged 32-row parts
constexpr int kAgSlots = 128;  // AG: flag words reserved per (source rank, 128-row block): K slabs x row slices

// Variants of the fused protocols (GemmParams::variant).
//   AG 0: unicast ring — push CTAs copy the shard to each peer with bulk copies (works without multicast)
//   AG 1: NVLS broadcast — push CTAs multimem.st the shard once, the switch replicates it; every rank pushes its
//         128-row blocks in the same order, so block j of ALL shards arrives together and the GEMM walks remote
//         tiles block-major
//   RS 0: scatter — the epilogue stores each partial tile into the owner's slot, fp32 slot reduction as a tail
//   RS 1: NVLS pull — the epilogue stores the partial tile LOCALLY and flags the owner; idle warps of the owner's
//         GEMM CTAs multimem.ld_reduce finished tiles while later tiles are computed (no scatter traffic, no tail);
//         every rank walks the owners round-robin so that all owners reduce all the time
enum GemmVariant : int { VARIANT_LEGACY = 0, VARIANT_NVLS = 1 };

struct GemmParams {
  CUtensorMap tma_a[2];     // AG: gathered workspace for parity 0/1; otherwise [0] only
  CUtensorMap tma_a_local;  // AG: this rank's own shard [M/W, K]
  CUtensorMap tma_b;
  SymmCtx ctx;
  void* c;        // PLAIN/AG: output [M, N]; RS: scatter buffer (symm, [2][W][M/W][N]); AR: symm tile buffer [2][M][ldc]
  void* out;      // RS: final output [M/W, N]; AR: final output [M, ldc] (symmetric when two-shot)
  long long ldc;  // row stride of c in elements
  int M, N, K;
  int mode;
  int is_bf16;
  int num_comm_ctas;  // CTAs [0, num_comm_ctas) are comm CTAs, the rest run the GEMM roles
  int group_m;        // L2 swizzle: m-blocks per group (PLAIN/AR)
  // distributed state
  const void* a_local;  // AG: own shard (rows_per_rank x K), plain device pointer
  void* a_ws;           // AG: symmetric gather workspace [2][M][K]
  uint32_t* flags;      // symmetric flag area (layout depends on mode)
  uint32_t* flags2;     // AR two-shot: "reduced tile landed" flags [tiles]
  uint32_t* counters;   // local tile counters (RS: per m-block)
  PhaseState ps;
  int rows_per_rank;  // M / W for AG/RS, M otherwise
  int ar_one_shot;    // AR: every rank reduces every tile (small outputs)
  int use_mc;         // AR: NVLS available
  int transpose_out;  // PLAIN: C is written transposed (swap-AB skinny GEMMs)
  int variant;             // GemmVariant
  int ag_slices;           // AG 1: push CTAs cooperating on one 128-row block (row slices), power of two <= 32
  int ag_kslabs;           // AG 1: a block is pushed, flagged and consumed in this many K slabs; ag_slices * ag_kslabs <= kAgSlots
  int ag_slab_kblocks;     // AG 1: 64-wide k blocks per slab
  int rs_unit_rows;        // RS 1: rows of a tile one reducer warp sums per unit (8, 16 or 32)
  int debug;               // ablations: 1 = RS reducers do nothing, 2 = no publish fence, 4 = reducers do not wait for flags (1, 2, 4: wrong results, timing only); 8 = RS publishes with a system-scope fence + release store
  long long a_lda;         // AG: row stride of a_local in elements
  size_t ws_parity_bytes;  // distance between the two parity halves of the workspace: a constant of the CONTEXT, so that
                           // calls of different sizes on one context never overlap the other parity's live data
  int flag_row;            // AG / RS (NVLS) / AR: flag words per SOURCE RANK. 0 = the row follows the call (bps * slots, bps * tiles_n, or tiles):
                           // the measured default. > 0 (TDIST_FIXED_FLAG_LAYOUT=1: flags_words / W, a constant of the
                           // context): every flag word has ONE writer rank over all calls, so calls of different M on one
                           // context cannot meet in a word under two meanings (profiles/protocol_sim.md)
  // grouped (MoE) GEMM: block_expert[m_blk] selects the expert whose [N, K] weight slab (rows
  // expert*N .. of B) multiplies the 128-row block; negative = padding block, skipped by all roles
  const int* block_expert;
};

template <int BLOCK_N, int STAGES>
struct GemmSmem {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiBytes = 4 * 32 * 128;  // 4 warps x 32 rows x 128 B
  static constexpr int kBarOffset = STAGES * kStageBytes + kEpiBytes;
  static constexpr int kGemmTotal = kBarOffset + (2 * STAGES + 4) * 8 + 16;
  static constexpr int kPushTotal = 6 * 32 * 1024 + 64;  // AG comm CTA ring
  static constexpr int kTotal = kGemmTotal > kPushTotal ? kGemmTotal : kPushTotal;
};

template <int BLOCK_N>
__host__ __device__ constexpr int tmem_cols() {
  return 2 * BLOCK_N <= 32 ? 32 : 2 * BLOCK_N <= 64 ? 64 : 2 * BLOCK_N <= 128 ? 128 : 2 * BLOCK_N <= 256 ? 256 : 512;
}

// ---- tile order -------------------------------------------------------------
struct TileCoord {
  int shard;    // AG: source rank of the A rows; RS: owner rank of the C rows; else 0
  int blk;      // 128-row block inside the shard
  int row0;     // first global row of the tile
  int row_end;  // exclusive global row limit for stores (end of shard / matrix)
  int n_blk;
};
__host__ __device__ __forceinline__ int blocks_per_shard(const GemmParams& p) { return (p.rows_per_rank + kBlockM - 1) / kBlockM; }

__host__ __device__ __forceinline__ TileCoord tile_coord(const GemmParams& p, int t, int tiles_m, int tiles_n) {
  TileCoord tc;
  if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
    // owner-interleaved, the same walk on every rank: tile t belongs to owner t % W, so all owners receive finished
    // tiles at the same steady rate and their reducers never idle
    const int W = p.ctx.world;
    const int bps = blocks_per_shard(p);
    tc.shard = t % W;
    const int i = t / W;
    tc.blk = i % bps;
    tc.n_blk = i / bps;
    tc.row0 = tc.shard * p.rows_per_rank + tc.blk * kBlockM;
    tc.row_end = (tc.shard + 1) * p.rows_per_rank;
  } else if (p.mode == GEMM_AG && p.variant == VARIANT_NVLS && t >= blocks_per_shard(p) * tiles_n) {
    // remote shards, block-major: (block 0 of rank+1, rank+2, ...), (block 1 of ...): the order in which the
    // broadcast pushes of all ranks land
    const int W = p.ctx.world;
    const int bps = blocks_per_shard(p);
    const int r = t - bps * tiles_n;
    const int unit = r / tiles_n;
    tc.n_blk = r - unit * tiles_n;
    tc.blk = unit / (W - 1);
    tc.shard = (p.ctx.rank + 1 + unit % (W - 1)) % W;
    tc.row0 = tc.shard * p.rows_per_rank + tc.blk * kBlockM;
    tc.row_end = (tc.shard + 1) * p.rows_per_rank;
  } else if (p.mode == GEMM_AG || p.mode == GEMM_RS) {
    // shard-major: AG consumes own shard first then rank+1, ...; RS produces the
    // segment of rank+1 first and its own last (same rotation as the reference,
    // allgather_gemm.py:224-229 / gemm_reduce_scatter.py:188,198).
    const int W = p.ctx.world;
    const int bps = blocks_per_shard(p);
    const int per_shard = bps * tiles_n;
    const int step = t / per_shard;
    const int r = t - step * per_shard;
    tc.shard = (p.ctx.rank + step + (p.mode == GEMM_RS ? 1 : 0)) % W;
    // inside a shard: group_m row blocks per column sweep, so that a B tile is reused group_m times from L2
    // (with few ranks a shard is many row blocks and B does not fit in L2)
    const int gm = p.group_m;
    const int per_group = gm * tiles_n;
    const int g = r / per_group;
    const int first_m = g * gm;
    const int rows = min(gm, bps - first_m);
    const int rr = r - g * per_group;
    tc.blk = first_m + rr % rows;
    tc.n_blk = rr / rows;
    tc.row0 = tc.shard * p.rows_per_rank + tc.blk * kBlockM;
    tc.row_end = (tc.shard + 1) * p.rows_per_rank;
  } else {
    // grouped ordering: walk group_m m-blocks per column sweep for L2 reuse of B
    const int gm = p.group_m;
    const int per_group = gm * tiles_n;
    const int g = t / per_group;
    const int first_m = g * gm;
    const int rows = min(gm, tiles_m - first_m);
    const int r = t - g * per_group;
    tc.shard = 0;
    tc.blk = first_m + r % rows;
    tc.n_blk = r / rows;
    tc.row0 = tc.blk * kBlockM;
    tc.row_end = p.M;
  }
  return tc;
}

// ---- comm CTA bodies ----------------------------------------------------------
// AG producer: push my shard to every peer, 128-row block by block, in the order
// peers will consume it (peer rank-1 needs it first). One thread drives a ring
// of bulk copies global -> smem -> peer global; the flag for (src=me, block) is
// released on the peer after the block's stores have completed.
__device__ __forceinline__ void ag_push_body(const GemmParams& p, uint8_t* smem, uint32_t epoch, int comm_idx,
                                             int num_comm) {
  constexpr int kChunk = 32 * 1024;
  constexpr int kSlots = 6;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSlots * kChunk);
  if (threadIdx.x == 0) {
    for (int i = 0; i < kSlots; ++i) ptx::mbar_init(bars + i, 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int W = p.ctx.world, rank = p.ctx.rank;
  const int bps = blocks_per_shard(p);
  const size_t row_bytes = size_t(p.K) * 2;
  const int parity = (epoch >> 1) & 1;
  const size_t ws_parity_off = size_t(parity) * p.ws_parity_bytes;
  const char* src_base = static_cast<const char*>(p.a_local);
  uint32_t bar_phase = 0;  // bit s = parity of slot s
  // work items: (step t = 1..W-1, block j, part q), flattened in consumption order and dealt round-robin to
  // the comm CTAs, so that at any time all of them work on the blocks the peers need next
  const int total = (W - 1) * bps * kAgParts;
  for (int item = comm_idx; item < total; item += num_comm) {
    const int t = 1 + item / (bps * kAgParts);
    const int j = (item / kAgParts) % bps;
    const int q = item % kAgParts;
    const int peer = (rank - t + W) % W;
    const int blk_rows = min(kBlockM, p.rows_per_rank - j * kBlockM);
    const int r0 = q * (kBlockM / kAgParts);
    const int rows = min(kBlockM / kAgParts, blk_rows - r0);
    if (rows > 0) {
      const size_t part_bytes = size_t(rows) * row_bytes;
      const char* src = src_base + (size_t(j) * kBlockM + r0) * row_bytes;  // (the ring needs a contiguous shard: lda == K)
      char* dst = symm_at(p.ctx, static_cast<char*>(p.a_ws), peer) + ws_parity_off +
                  (size_t(rank) * p.rows_per_rank + size_t(j) * kBlockM + r0) * row_bytes;
      const int n_chunks = int((part_bytes + kChunk - 1) / kChunk);
      int issued = 0;
      for (; issued < n_chunks && issued < kSlots; ++issued) {
        const uint32_t bytes = uint32_t(min(size_t(kChunk), part_bytes - size_t(issued) * kChunk));
        ptx::mbar_arrive_expect_tx(bars + issued, bytes);
        ptx::bulk_g2s(smem + issued * kChunk, src + size_t(issued) * kChunk, bytes, bars + issued);
      }
      for (int k = 0; k < n_chunks; ++k) {
        const int slot = k % kSlots;
        const uint32_t bytes = uint32_t(min(size_t(kChunk), part_bytes - size_t(k) * kChunk));
        ptx::mbar_wait(bars + slot, (bar_phase >> slot) & 1);
        bar_phase ^= 1u << slot;
        ptx::bulk_s2g(dst + size_t(k) * kChunk, smem + slot * kChunk, bytes);
        ptx::bulk_commit();
        // refill the slot used by the previous chunk once its smem has been read
        if (k >= 1 && issued < n_chunks) {
          ptx::bulk_wait_read<1>();
          const int rs = (k - 1) % kSlots;
          const uint32_t b2 = uint32_t(min(size_t(kChunk), part_bytes - size_t(issued) * kChunk));
          ptx::mbar_arrive_expect_tx(bars + rs, b2);
          ptx::bulk_g2s(smem + rs * kChunk, src + size_t(issued) * kChunk, b2, bars + rs);
          ++issued;
        }
      }
      ptx::bulk_wait<0>();  // all stores of this part are complete
      ptx::fence_proxy_async();
      ptx::fence_acq_rel_sys();
    }
    // flags layout (AG): [W src][blocks_per_shard (or the fixed row, GemmParams::flag_row)][kAgParts]; empty parts are flagged too
    notify(p.ctx, p.flags + size_t(rank) * (p.flag_row > 0 ? p.flag_row : bps * kAgParts) + j * kAgParts + q, peer, epoch);
  }
}

// AG producer, NVLS variant: the push CTAs broadcast this rank's shard with multimem.st — every 16 bytes leave the GPU
// once and the switch delivers them to all ranks' workspaces. An item is (128-row block j, K slab s, row slice sl):
// `ag_slices` CTAs split the rows of a block, and a block may be sent in K slabs so that the consumers' K loops can
// follow the arrival. Items are dealt round-robin to the push CTAs in (block, slab, slice) order. All threads of the
// CTA move data (8 independent 16-byte loads per thread in flight before the first store), then the CTA meets, one
// thread makes the stores visible system-wide and releases the item's flag on every rank.
// Measured alternative (profiles/README.md): a publisher warp that fences while the other warps keep streaming is
// SLOWER (0.173 vs 0.145 ms at 8 ranks) — the system fence then also waits for the younger stores, so flags come late.
__device__ __forceinline__ void ag_push_mc_body(const GemmParams& p, uint32_t epoch, int comm_idx, int num_comm) {
  const int rank = p.ctx.rank;
  const int bps = blocks_per_shard(p);
  const int slices = p.ag_slices, slabs = p.ag_kslabs;
  const int rows_per_slice = kBlockM / slices;
  const int slab_cols = p.ag_slab_kblocks * kBlockK;
  const int parity = (epoch >> 1) & 1;
  const char* src_base = static_cast<const char*>(p.a_local);
  char* dst_base = mc_ptr(p.ctx, static_cast<char*>(p.a_ws)) + size_t(parity) * p.ws_parity_bytes +
                   size_t(rank) * p.rows_per_rank * size_t(p.K) * 2;
  const int items = bps * slabs * slices;
  constexpr int kBatch = 8;
  for (int item = comm_idx; item < items; item += num_comm) {
    const int sl = item % slices;
    const int js = item / slices;
    const int j = js / slabs, sb = js - j * slabs;
    const int r0 = j * kBlockM + sl * rows_per_slice;
    const int rows = max(0, min(rows_per_slice, p.rows_per_rank - r0));
    const int col0 = sb * slab_cols;
    const int seg_vecs = max(0, min(slab_cols, p.K - col0)) / 8;  // 16-byte vectors per row of this slab
    const int n_vec = rows * seg_vecs;
    for (int i0 = threadIdx.x; i0 < n_vec; i0 += blockDim.x * kBatch) {
      int4 v[kBatch];
      size_t doff[kBatch];
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const int i = i0 + b * blockDim.x;
        if (i < n_vec) {
          const int r = i / seg_vecs, cv = i - r * seg_vecs;
          v[b] = ptx::ld_nc_v4(src_base + (size_t(r0 + r) * p.a_lda + col0) * 2 + size_t(cv) * 16);
          doff[b] = (size_t(r0 + r) * p.K + col0) * 2 + size_t(cv) * 16;
        }
      }
#pragma unroll
      for (int b = 0; b < kBatch; ++b)
        if (i0 + b * blockDim.x < n_vec) ptx::multimem_st_v4(dst_base + doff[b], v[b]);
    }
    __syncthreads();
    // flags layout (AG, NVLS): [W src][blocks_per_shard][kAgSlots], slot = slab * slices + slice
    if (threadIdx.x == 0) {
      ptx::fence_acq_rel_sys();
      ptx::multimem_st_b32(mc_ptr(p.ctx, p.flags + size_t(rank) * (p.flag_row > 0 ? p.flag_row : bps * kAgSlots) + j * kAgSlots + sb * slices + sl), epoch);
    }
  }
}

// RS consumer, NVLS variant: run by the two otherwise idle warps of every GEMM CTA. A unit is rs_unit_rows rows of one
// (128-row block, n block) tile of MY segment; units are dealt round-robin to the reducer warps of the grid in the
// order the tiles are produced (the same on every rank). The warp waits until all W ranks have flagged the tile, sums
// the W partial copies in the switch with multimem.ld_reduce (fp32 accumulation) and writes the result to `out`.
// Without a multicast mapping the same loop sums plain peer loads.
template <typename Tag, int BLOCK_N>
__device__ __forceinline__ void rs_pull_body(const GemmParams& p, uint32_t epoch, int cw, int ncw, int tiles_n, bool pair_order) {
  const int W = p.ctx.world;
  const int lane = threadIdx.x & 31;
  const int bps = blocks_per_shard(p);
  const int bps2 = (p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM);
  const int parity = (epoch >> 1) & 1;
  const int unit_rows = p.rs_unit_rows;
  const int kSub = kBlockM / unit_rows;
  const int seg_tiles = (pair_order ? bps2 * 2 : bps) * tiles_n;  // 128-row tiles of my segment in production order
  const int units = seg_tiles * kSub;
  const char* part = static_cast<const char*>(p.c) + size_t(parity) * p.ws_parity_bytes;
  const char* part_mc = mc_ptr(p.ctx, part);
  char* out = static_cast<char*>(p.out);
  const bool use_mc = p.ctx.mc_base != nullptr && p.use_mc;
  constexpr int kBatch = 8;
  if (p.debug & 1) return;
  for (int u = cw; u < units; u += ncw) {
    const int ti = u / kSub, sc = u - ti * kSub;
    int blk, n_blk;
    if (pair_order) {
      // pair tiles are produced as (pair block i % bps2, n block i / bps2); both 128-row halves at once
      const int i = ti >> 1;
      blk = (i % bps2) * 2 + (ti & 1);
      n_blk = i / bps2;
    } else {
      blk = ti % bps;
      n_blk = ti / bps;
    }
    const int r0 = blk * kBlockM + sc * unit_rows;
    if (blk >= bps || r0 >= p.rows_per_rank) continue;
    // flags layout (RS, NVLS): [W src][bps][tiles_n], written by src when its partial tile is stored
    if (lane < W && !(p.debug & 4))
      wait_ge_sys_poll_relaxed(p.flags + size_t(lane) * (p.flag_row > 0 ? p.flag_row : bps * tiles_n) + blk * tiles_n + n_blk, epoch);
    __syncwarp();
    const int rows = min(unit_rows, p.rows_per_rank - r0);
    const int col0 = n_blk * BLOCK_N;
    const int vpr = min(BLOCK_N, p.N - col0) / 8;  // 16-byte vectors per row of this tile
    const int n_vec = rows * vpr;
    for (int i0 = lane; i0 < n_vec; i0 += 32 * kBatch) {
      int4 v[kBatch];
      size_t ooff[kBatch];
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const int i = i0 + b * 32;
        if (i < n_vec) {
          const int r = i / vpr, cv = i - r * vpr;
          ooff[b] = (size_t(r0 + r) * p.N + col0) * 2 + size_t(cv) * 16;
          const size_t poff = (size_t(p.ctx.rank) * p.rows_per_rank * size_t(p.N)) * 2 + ooff[b];
          if (use_mc) {
            v[b] = Pack<Tag>::mc_ld_reduce(part_mc + poff);
          } else {
            float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 1
            for (int s2 = 0; s2 < W; ++s2) Pack<Tag>::add(acc, ptx::ld_v4(symm_at(p.ctx, part, (p.ctx.rank + s2) % W) + poff));
            v[b] = Pack<Tag>::cvt(acc);
          }
        }
      }
#pragma unroll
      for (int b = 0; b < kBatch; ++b)
        if (i0 + b * 32 < n_vec) *reinterpret_cast<int4*>(out + ooff[b]) = v[b];
    }
  }
}

// RS tail: out[r, :] = sum_src scatter[parity][src][r, :] for my segment,
// gated per 128-row block by the W arrival flags. Executed by every CTA of the
// grid after its GEMM work, so the tail runs at full HBM bandwidth.
template <typename Tag>
__device__ __forceinline__ void rs_reduce_body(const GemmParams& p, uint32_t epoch) {
  // Work unit = (128-row block j of my segment, 128-column slice): units are independent, so the tail is not a chain
  // of (flag wait -> barrier -> one load round trip) per block any more (measured: ~3 us per block, 32 blocks at
  // TP2). Every thread keeps W x 4 sixteen-byte loads in flight.
  const int W = p.ctx.world;
  const int bps = blocks_per_shard(p);
  const int parity = (epoch >> 1) & 1;
  const size_t seg_elems = size_t(p.rows_per_rank) * p.N;
  const int4* scat = reinterpret_cast<const int4*>(static_cast<const char*>(p.c) + size_t(parity) * p.ws_parity_bytes);
  int4* out = static_cast<int4*>(p.out);
  const size_t seg_vecs = seg_elems / 8;
  const int row_vecs = p.N / 8;                 // 16-byte vectors per row
  constexpr int kSliceVecs = 16;                // 128 columns
  const int slices = (row_vecs + kSliceVecs - 1) / kSliceVecs;
  const int units = bps * slices;
  int last_j = -1;
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int j = u / slices, sl = u - j * slices;
    if (j != last_j) {
      // flags layout (RS): [W src][bps (or the fixed row, GemmParams::flag_row)], written by src when its block j of my segment is complete
      if (threadIdx.x < W) wait_ge_sys(p.flags + size_t(threadIdx.x) * (p.flag_row > 0 ? p.flag_row : bps) + j, epoch);
      __syncthreads();
      last_j = j;
    }
    const int rows = min(kBlockM, p.rows_per_rank - j * kBlockM);
    const int v0 = sl * kSliceVecs, nv = min(kSliceVecs, row_vecs - v0);
    const int total = rows * nv;
    constexpr int kBatch = 4;
    for (int i0 = threadIdx.x; i0 < total; i0 += blockDim.x * kBatch) {
      size_t off[kBatch];
      float acc[kBatch][8];
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const int i = i0 + b * blockDim.x;
        const int r = i / nv, c = i - r * nv;
        off[b] = (size_t(j) * kBlockM + r) * row_vecs + v0 + c;
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[b][e] = 0.f;
      }
#pragma unroll 2
      for (int sidx = 0; sidx < W; ++sidx) {
        int4 v[kBatch];
#pragma unroll
        for (int b = 0; b < kBatch; ++b)
          if (i0 + b * blockDim.x < total) v[b] = ptx::ld_v4(scat + size_t(sidx) * seg_vecs + off[b]);
#pragma unroll
        for (int b = 0; b < kBatch; ++b)
          if (i0 + b * blockDim.x < total) Pack<Tag>::add(acc[b], v[b]);
      }
#pragma unroll
      for (int b = 0; b < kBatch; ++b)
        if (i0 + b * blockDim.x < total) out[off[b]] = Pack<Tag>::cvt(acc[b]);
    }
  }
}

// AR consumer: run by warp 2 of every GEMM CTA (no SM is reserved for communication). Tiles go to a symmetric buffer;
// warp 3 of the producing CTA flags them (GPU-scope fence + plain system store, see the RS publisher). A unit is
// rs_unit_rows rows of one tile, dealt round-robin to the reducer warps of the grid in production order.
//   one-shot (decode sizes): every rank sums every tile itself (flags are broadcast to all ranks);
//   two-shot (prefill sizes): rank r sums the tiles t with t % W == r and broadcasts the result with multimem.st; a
//   closing exchange of one "all my tiles are out" word per rank makes `out` complete on every rank at kernel end.
// flags (AR): [W src][tiles] on each rank; flags2: [W] closing words.
template <typename Tag, int BLOCK_N>
__device__ __forceinline__ void ar_reduce_body(const GemmParams& p, uint32_t epoch, int cw, int ncw, int tiles_m, int tiles_n) {
  const int W = p.ctx.world;
  const int lane = threadIdx.x & 31;
  const int tiles = tiles_m * tiles_n;
  const int parity = (epoch >> 1) & 1;
  const int unit_rows = p.rs_unit_rows;
  const int kSub = kBlockM / unit_rows;
  char* sym_c = static_cast<char*>(p.c) + size_t(parity) * p.ws_parity_bytes;
  char* out = static_cast<char*>(p.out);
  const int owned = p.ar_one_shot ? tiles : (tiles - p.ctx.rank + W - 1) / W;
  const int units = owned * kSub;
  constexpr int kBatch = 8;
  for (int u = cw; u < units; u += ncw) {
    const int k = u / kSub, sc = u - k * kSub;
    const int t = p.ar_one_shot ? k : p.ctx.rank + k * W;
    const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
    const int r0 = tc.row0 + sc * unit_rows;
    if (r0 >= p.M) continue;
    if (lane < W) wait_ge_sys_poll_relaxed(p.flags + size_t(lane) * (p.flag_row > 0 ? p.flag_row : tiles) + t, epoch);
    __syncwarp();
    const int rows = min(unit_rows, p.M - r0);
    const int col0 = tc.n_blk * BLOCK_N;
    const int vpr = min(BLOCK_N, p.N - col0) / 8;
    const int n_vec = rows * vpr;
    for (int i0 = lane; i0 < n_vec; i0 += 32 * kBatch) {
      int4 v[kBatch];
      size_t off[kBatch];
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const int i = i0 + b * 32;
        if (i < n_vec) {
          const int r = i / vpr, cv = i - r * vpr;
          off[b] = (size_t(r0 + r) * p.ldc + col0) * 2 + size_t(cv) * 16;
          if (p.use_mc) {
            v[b] = Pack<Tag>::mc_ld_reduce(mc_ptr(p.ctx, sym_c) + off[b]);
          } else {
            float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 1
            for (int s2 = 0; s2 < W; ++s2) Pack<Tag>::add(acc, ptx::ld_v4(symm_at(p.ctx, sym_c, (p.ctx.rank + s2) % W) + off[b]));
            v[b] = Pack<Tag>::cvt(acc);
          }
        }
      }
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        if (i0 + b * 32 >= n_vec) continue;
        if (p.ar_one_shot) {
          *reinterpret_cast<int4*>(out + off[b]) = v[b];
        } else if (p.use_mc) {
          ptx::multimem_st_v4(mc_ptr(p.ctx, out) + off[b], v[b]);
        } else {
#pragma unroll 1
          for (int s2 = 0; s2 < W; ++s2) ptx::st_v4(symm_at(p.ctx, out, (p.ctx.rank + s2) % W) + off[b], v[b]);
        }
      }
    }
  }
  if (!p.ar_one_shot) {
    // closing exchange: the last reducer warp of this rank to finish tells every rank that all my tiles are out
    __syncwarp();
    if (lane == 0) {
      ptx::fence_acq_rel_sys();
      if (ptx::atom_add_acq_rel_gpu(p.counters, 1) == static_cast<uint32_t>(ncw - 1)) {
        *p.counters = 0;
        ptx::fence_acq_rel_sys();
        for (int s2 = 0; s2 < W; ++s2) ptx::st_relaxed_sys(symm_at(p.ctx, p.flags2 + p.ctx.rank, (p.ctx.rank + s2) % W), epoch);
      }
    }
    __syncwarp();
    if (lane < W) wait_ge_sys_poll_relaxed(p.flags2 + lane, epoch);
    __syncwarp();
  }
}

// ---- the kernel ---------------------------------------------------------------
template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_sm100_kernel(const __grid_constant__ GemmParams p) {
  using L = GemmSmem<BLOCK_N, STAGES>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int tiles_m = (p.mode == GEMM_AG || p.mode == GEMM_RS) ? blocks_per_shard(p) * p.ctx.world
                                                               : (p.M + kBlockM - 1) / kBlockM;
  const int tiles = tiles_m * tiles_n;
  const int num_k = (p.K + kBlockK - 1) / kBlockK;
  const int num_gemm = gridDim.x - p.num_comm_ctas;
  const int gemm_cta = static_cast<int>(blockIdx.x) - p.num_comm_ctas;

  uint32_t epoch = 0, phase_base = 0;
  if (p.mode != GEMM_PLAIN) {
    phase_base = phase_begin(p.ps);
    epoch = phase_base + 2;  // flags of this call carry `epoch`; parity = (epoch >> 1) & 1
  }
  const int parity = (epoch >> 1) & 1;

  if (gemm_cta < 0) {
    // ------------------------------------------------------------ comm CTA
    if (p.mode == GEMM_AG) {
      if (p.variant == VARIANT_NVLS)
        ag_push_mc_body(p, epoch, blockIdx.x, p.num_comm_ctas);
      else
        ag_push_body(p, smem, epoch, blockIdx.x, p.num_comm_ctas);
    }
  } else {
    // ------------------------------------------------------------ GEMM CTA
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
        ptx::mbar_init(full_bar + i, 1);
        ptx::mbar_init(empty_bar + i, 1);
      }
      for (int i = 0; i < 2; ++i) {
        ptx::mbar_init(tmem_full + i, 1);
        ptx::mbar_init(tmem_empty + i, 4);  // one arrival per epilogue warp
      }
      *stored_count = 0;
      ptx::fence_barrier_init();
    }
    if (warp == 2) ptx::tmem_alloc<kTmemCols>(tmem_ptr_smem);
    ptx::tcgen05_fence_before();
    __syncthreads();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
      // ======================================================== TMA producer
      // The whole warp walks the tiles (so that its lanes can poll the arrival flags of a block in parallel); lane 0
      // issues the copies.
      {
        int stage = 0;
        uint32_t ph = 0;
        const int bps = blocks_per_shard(p);
        for (int t = gemm_cta; t < tiles; t += num_gemm) {
          const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
          int b_row = tc.n_blk * BLOCK_N;
          if (p.block_expert != nullptr) {
            const int e = p.block_expert[tc.blk];
            if (e < 0) continue;
            b_row += e * p.N;
          }
          // [sim:ag_wait begin] (csrc/sim runs the text between these markers against ag_push_mc_body on the CPU)
          const CUtensorMap* ta = &p.tma_a[0];
          int a_row = tc.row0;
          const uint32_t* slab_flags = nullptr;
          if (p.mode == GEMM_AG) {
            if (tc.shard == p.ctx.rank) {
              ta = &p.tma_a_local;
              a_row = tc.blk * kBlockM;
            } else {
              ta = &p.tma_a[parity];
              if (p.variant != VARIANT_NVLS) {
                // wait until every part of block (src, blk) has landed in my workspace
                if (lane < kAgParts)
                  wait_ge_sys_poll_relaxed(p.flags + size_t(tc.shard) * (p.flag_row > 0 ? p.flag_row : bps * kAgParts) + tc.blk * kAgParts + lane, epoch);
                __syncwarp();
                ptx::fence_proxy_async();
              } else {
                slab_flags = p.flags + size_t(tc.shard) * (p.flag_row > 0 ? p.flag_row : bps * kAgSlots) + tc.blk * kAgSlots;
              }
            }
          }
          for (int kb = 0; kb < num_k; ++kb) {
            if (slab_flags != nullptr && kb % p.ag_slab_kblocks == 0) {
              // NVLS push: the K loop follows the arrival of the block's K slabs (lanes poll the row slices of a slab)
              if (lane < p.ag_slices) wait_ge_sys_poll_relaxed(slab_flags + (kb / p.ag_slab_kblocks) * p.ag_slices + lane, epoch);
              __syncwarp();
              ptx::fence_proxy_async();
            }
            // [sim:ag_wait end]
            if (lane == 0) {
              ptx::mbar_wait(empty_bar + stage, ph ^ 1);
              ptx::mbar_arrive_expect_tx(full_bar + stage, L::kStageBytes);
              ptx::tma_load_2d(smem_a + stage * L::kABytes, ta, full_bar + stage, kb * kBlockK, a_row);
              ptx::tma_load_2d(smem_b + stage * L::kBBytes, &p.tma_b, full_bar + stage, kb * kBlockK, b_row);
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
      // ======================================================== MMA issuer
      if (lane == 0) {
        const uint32_t idesc = ptx::make_idesc_f16(kBlockM, BLOCK_N, p.is_bf16 != 0);
        int stage = 0;
        uint32_t ph = 0;
        int acc = 0;
        uint32_t acc_ph = 0;
        for (int t = gemm_cta; t < tiles; t += num_gemm) {
          if (p.block_expert != nullptr && p.block_expert[tile_coord(p, t, tiles_m, tiles_n).blk] < 0) continue;
          ptx::mbar_wait(tmem_empty + acc, acc_ph ^ 1);
          ptx::tcgen05_fence_after();
          const uint32_t d_tmem = tmem_base + acc * kAccStride;
          for (int kb = 0; kb < num_k; ++kb) {
            ptx::mbar_wait(full_bar + stage, ph);
            ptx::tcgen05_fence_after();
            const uint64_t da = ptx::make_smem_desc_sw128(smem_a + stage * L::kABytes);
            const uint64_t db = ptx::make_smem_desc_sw128(smem_b + stage * L::kBBytes);
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k) {
              // +32 bytes per K step inside the 128-byte swizzle row (encoded >> 4)
              ptx::umma_f16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            }
            ptx::umma_commit(empty_bar + stage);  // frees the smem stage when these MMAs retire
            if (++stage == STAGES) {
              stage = 0;
              ph ^= 1;
            }
          }
          ptx::umma_commit(tmem_full + acc);  // accumulator complete -> epilogue
          if (++acc == 2) {
            acc = 0;
            acc_ph ^= 1;
          }
        }
      }
    } else if (warp >= 4) {
      // ======================================================== epilogue
      const int q = warp - 4;  // TMEM lane quarter this warp may access (warp id % 4)
      uint8_t* stg = smem_epi + q * (32 * 128);
      int acc = 0;
      uint32_t acc_ph = 0;
      const int W = p.ctx.world;
      // RS/AR: a finished tile is published (system fence + counter/flag) one tile LATER, right before the next
      // tile's stores are issued: by then its NVLink stores have drained, so the fence costs nothing instead of
      // stalling the epilogue for a link round trip per tile.
      int pend_t = -1, pend_shard = 0, pend_blk = 0, pend_nblk = 0;
      // [sim:epi_publish begin] (csrc/sim runs this lambda against rs_reduce_body on the CPU)
      auto publish = [&](int t_done, int shard, int blk, int n_blk) {
        if (!(p.debug & 2)) ptx::fence_acq_rel_sys();
        if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
          // my partial tile is in my own memory: tell the owner it may be summed
          const int bps = blocks_per_shard(p);
          notify(p.ctx, p.flags + (size_t(p.ctx.rank) * bps + blk) * tiles_n + n_blk, shard, epoch);
        } else if (p.mode == GEMM_RS) {
          const int bps = blocks_per_shard(p);
          const int cidx = shard * bps + blk;
          const uint32_t done = ptx::atom_add_acq_rel_gpu(p.counters + cidx, 1) + 1;
          if (done == static_cast<uint32_t>(tiles_n)) {
            p.counters[cidx] = 0;
            ptx::fence_acq_rel_sys();
            notify(p.ctx, p.flags + size_t(p.ctx.rank) * (p.flag_row > 0 ? p.flag_row : bps) + blk, shard, epoch);
          }
        } else {
          ptx::st_release_sys(p.flags + t_done, epoch);
        }
      };
      // [sim:epi_publish end]
      for (int t = gemm_cta; t < tiles; t += num_gemm) {
        const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
        if (p.block_expert != nullptr && p.block_expert[tc.blk] < 0) continue;
        ptx::mbar_wait(tmem_full + acc, acc_ph);
        ptx::tcgen05_fence_after();
        if (pend_t >= 0 && warp == 4 && lane == 0) publish(pend_t, pend_shard, pend_blk, pend_nblk);
        pend_t = -1;
        // destination of this tile
        char* dst_base = static_cast<char*>(p.c);
        long long ld = p.ldc;
        int row0 = tc.row0 + q * 32;  // first global row handled by this warp
        int row_limit = tc.row_end;
        if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
          // partial tile stays local, in the [M, N] layout that is identical (and multicast-mapped) on every rank
          dst_base += size_t(parity) * p.ws_parity_bytes;
          ld = p.N;
        } else if (p.mode == GEMM_RS) {
          const size_t seg_elems = size_t(p.rows_per_rank) * p.N;
          dst_base = symm_at(p.ctx, static_cast<char*>(p.c), tc.shard) + size_t(parity) * p.ws_parity_bytes +
                     size_t(p.ctx.rank) * seg_elems * 2;
          row0 -= tc.shard * p.rows_per_rank;
          row_limit = p.rows_per_rank;
          ld = p.N;
        } else if (p.mode == GEMM_AR) {
          dst_base += size_t(parity) * p.ws_parity_bytes;
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
            // last TMEM read of this tile: hand the accumulator back to the MMA warp
            ptx::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(tmem_empty + acc);
          }
          if (!p.transpose_out) {
            // pack to 16-bit, write my row (128 B) into the staging tile with a 16-byte XOR swizzle
            uint32_t pk[32];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float a0 = __uint_as_float(v0[2 * i]), a1 = __uint_as_float(v0[2 * i + 1]);
              const float b0 = __uint_as_float(v1[2 * i]), b1 = __uint_as_float(v1[2 * i + 1]);
              pk[i] = p.is_bf16 ? ptx::pack_bf16x2(a0, a1) : ptx::pack_f16x2(a0, a1);
              pk[16 + i] = p.is_bf16 ? ptx::pack_bf16x2(b0, b1) : ptx::pack_f16x2(b0, b1);
            }
            __syncwarp();  // previous chunk's readers are done with the staging tile
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              int4 w = make_int4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
              *reinterpret_cast<int4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = w;
            }
            __syncwarp();
            // coalesced stores: 8 lanes cover one 128-byte row, 4 rows per instruction
            const int cj = lane & 7;
            const int gcol = col0 + c + cj * 8;
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int r = it * 4 + (lane >> 3);
              const int4 w = *reinterpret_cast<const int4*>(stg + r * 128 + ((cj ^ (r & 7)) << 4));
              const int grow = row0 + r;
              if (grow < row_limit && gcol < p.N && c + cj * 8 < BLOCK_N)
                ptx::st_v4(dst_base + (size_t(grow) * ld + gcol) * 2, w);
            }
          } else {
            // swap-AB skinny GEMM: accumulator row = output column. C[col][row] scalar stores,
            // a warp writes 32 consecutive output columns of one output row per instruction.
            const int grow = row0 + lane;
#pragma unroll
            for (int i = 0; i < 64; ++i) {
              const int gc = col0 + c + i;
              if (c + i < BLOCK_N && gc < p.N && grow < row_limit) {
                const float f = __uint_as_float(i < 32 ? v0[i & 31] : v1[i & 31]);
                if (p.is_bf16)
                  reinterpret_cast<__nv_bfloat16*>(dst_base)[size_t(gc) * ld + grow] = __float2bfloat16(f);
                else
                  reinterpret_cast<__half*>(dst_base)[size_t(gc) * ld + grow] = __float2half(f);
              }
            }
          }
        }
        if ((p.mode == GEMM_RS && p.variant == VARIANT_NVLS) || p.mode == GEMM_AR) {
          // The partial tile is in local memory. Telling the owner needs a system-scope fence, and that fence also
          // waits for whatever else this SM has in flight (measured: 24 us of the 8-rank op when the epilogue issued
          // it while the reducer warp's NVLS loads were outstanding). So the epilogue only counts the tile; warp 3
          // publishes it.
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (warp == 4 && lane == 0) {
            __threadfence_block();
            atomicAdd(stored_count, 1u);
          }
        } else if (p.mode == GEMM_RS) {
          // all 4 epilogue warps have issued their stores; remember the tile, publish it one tile later
          asm volatile("bar.sync 1, 128;" ::: "memory");
          pend_t = t;
          pend_shard = tc.shard;
          pend_blk = tc.blk;
          pend_nblk = tc.n_blk;
        }
        if (++acc == 2) {
          acc = 0;
          acc_ph ^= 1;
        }
      }
      if (pend_t >= 0 && warp == 4 && lane == 0) publish(pend_t, pend_shard, pend_blk, pend_nblk);
    } else if (p.mode == GEMM_RS && p.variant == VARIANT_NVLS) {
      if (warp == 3) {
        // ====================================================== warp 3: publishes this CTA's finished tiles
        if (lane == 0) {
          // [sim:rs_publish begin] (csrc/sim runs the text between these markers against rs_pull_body on the CPU)
          const int bps = blocks_per_shard(p);
          uint32_t k = 0;
          for (int t = gemm_cta; t < tiles; t += num_gemm, ++k) {
            const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
            while (*reinterpret_cast<volatile uint32_t*>(stored_count) <= k) {
            }
            __threadfence_block();
            uint32_t* flag = symm_at(
                p.ctx, p.flags + size_t(p.ctx.rank) * (p.flag_row > 0 ? p.flag_row : bps * tiles_n) + tc.blk * tiles_n + tc.n_blk, tc.shard);
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
          // [sim:rs_publish end]
        }
      } else {
        // ====================================================== warp 2: NVLS reducer of my own segment
        if (p.is_bf16)
          rs_pull_body<BF16, BLOCK_N>(p, epoch, gemm_cta, num_gemm, tiles_n, false);
        else
          rs_pull_body<F16, BLOCK_N>(p, epoch, gemm_cta, num_gemm, tiles_n, false);
      }
    }
    else if (p.mode == GEMM_AR) {
      if (warp == 3) {
        // ====================================================== warp 3: publishes this CTA's finished tiles
        if (lane == 0) {
          // [sim:ar_publish begin]
          uint32_t k = 0;
          const int W = p.ctx.world;
          for (int t = gemm_cta; t < tiles; t += num_gemm, ++k) {
            while (*reinterpret_cast<volatile uint32_t*>(stored_count) <= k) {
            }
            __threadfence_block();
            ptx::fence_acq_rel_gpu();  // the tile is in this GPU's memory: see the RS publisher
            uint32_t* flag = p.flags + size_t(p.ctx.rank) * (p.flag_row > 0 ? p.flag_row : tiles) + t;
            if (!p.ar_one_shot) {
              ptx::st_relaxed_sys(symm_at(p.ctx, flag, t % W), epoch);
            } else if (p.use_mc) {
              ptx::multimem_st_b32(mc_ptr(p.ctx, flag), epoch);
            } else {
              for (int s2 = 0; s2 < W; ++s2) ptx::st_relaxed_sys(symm_at(p.ctx, flag, (p.ctx.rank + s2) % W), epoch);
            }
          }
          // [sim:ar_publish end]
        }
      } else {
        // ====================================================== warp 2: cross-rank reducer
        if (p.is_bf16)
          ar_reduce_body<BF16, BLOCK_N>(p, epoch, gemm_cta, num_gemm, tiles_m, tiles_n);
        else
          ar_reduce_body<F16, BLOCK_N>(p, epoch, gemm_cta, num_gemm, tiles_m, tiles_n);
      }
    }
    // teardown of the GEMM roles
    ptx::tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<kTmemCols>(tmem_base);
  }

  // ------------------------------------------------------------ fused tails (all CTAs)
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
