// overlap_tutorial.cu — the two overlap protocols of this library, written out at tutorial size.
// (Counterpart of the reference's tutorials/07-overlapping-allgather-gemm.py and 08-overlapping-gemm-reduce-scatter.py,
// which build the same ideas from Triton primitives; here the primitives are the ones of csrc/device/symm.cuh.)
//
// The math is a plain CUDA-core tile product so that nothing distracts from the communication: the production kernels
// (csrc/kernels/gemm_sm100.cuh, gemm_sm100_2cta.cuh) have exactly this structure with TMA + tcgen05 in place of the
// inner loops. tutorials/07-*.py and 08-*.py run these kernels, check them against torch and explain each step.
#include "../kernels/common.cuh"

namespace tdist {
namespace tut {

constexpr int kTile = 64;      // rows of A per step of the protocol ("block")
constexpr int kThreads = 256;  // 16 x 16 threads, each owns a 4 x 4 patch of a 64 x 64 output tile

// C[rows r0.., cols c0..] (64 x 64) = A[r0.., :K] * B[c0.., :K]^T, fp32 accumulate, bf16 in / out. A and B are row-major
// with leading dimensions lda / ldb; rows >= a_rows and columns >= b_rows are skipped.
__device__ void tile_gemm(const __nv_bfloat16* A, long long lda, int a_rows, const __nv_bfloat16* B, long long ldb, int b_rows,
                          int K, int r0, int c0, float (&acc)[4][4]) {
  __shared__ float sa[kTile][17], sb[kTile][17];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int idx = threadIdx.x; idx < kTile * 16; idx += kThreads) {
      const int r = idx >> 4, kk = idx & 15;
      sa[r][kk] = (r0 + r < a_rows && k0 + kk < K) ? __bfloat162float(A[(long long)(r0 + r) * lda + k0 + kk]) : 0.f;
      sb[r][kk] = (c0 + r < b_rows && k0 + kk < K) ? __bfloat162float(B[(long long)(c0 + r) * ldb + k0 + kk]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(sa[ty * 4 + i][kk], sb[tx * 4 + j][kk], acc[i][j]);
    __syncthreads();
  }
}

// ================================================================= tutorial 7: AllGather + GEMM in ONE kernel
// Every rank owns `rows` rows of A. C[W * rows, N] = all_gather(A) * B^T.
//   * CTAs [0, push_ctas) are producers: they broadcast this rank's shard, one 64-row block at a time, into the gather
//     workspace of EVERY rank (multimem.st when the heap has a multicast mapping: the bytes leave this GPU once and
//     the switch replicates them; otherwise one plain store per peer) and then release the block's flag everywhere.
//   * all other CTAs are consumers: they walk the output tiles — own shard first (no waiting: it is read from `a`),
//     then block 0 of every remote shard, block 1, ... = the order in which the blocks arrive — and before touching a
//     remote block they spin on ITS flag only.
// Flags carry a monotone `epoch` (never reset) and the workspace has two halves used alternately, so calls can follow
// each other without any barrier between them. The halves are `parity_elems` apart — HALF THE WORKSPACE, a constant of the
// buffer and not a size of this call: with a stride of W * rows * K a call with fewer rows would put its half inside the
// region a slower rank is still reading for the previous, larger call (csrc/sim/protocol_sim.cc finds exactly that race).
// The flag of (source rank, block) sits at [rank * flag_blocks + block] with flag_blocks a constant of the flag array as
// well: a layout that moved with `rows` would let two ranks that are one call apart write the same word under different
// meanings — a straggler's older epoch could then overwrite a newer one and a waiter would never wake up (the simulator
// shows both the race and the hang for the reduce-scatter kernel below when the layout follows the call).
// ws: [2][parity_elems] symmetric, parity_elems >= W * rows * K; flags: [W][flag_blocks] symmetric, zero initialised once.
__global__ void __launch_bounds__(kThreads) tut_ag_gemm_kernel(SymmCtx ctx, const __nv_bfloat16* a, const __nv_bfloat16* b,
                                                               __nv_bfloat16* c, __nv_bfloat16* ws, size_t parity_elems,
                                                               uint32_t* flags, int flag_blocks, int rows, int N, int K,
                                                               int push_ctas, PhaseState ps) {
  const uint32_t base = phase_begin(ps);  // the same value in every CTA of this launch
  const uint32_t epoch = base + 2;
  const int parity = (epoch >> 1) & 1;
  const int W = ctx.world, rank = ctx.rank;
  const int blocks = (rows + kTile - 1) / kTile;
  __nv_bfloat16* my_ws = ws + size_t(parity) * parity_elems;
  if (int(blockIdx.x) < push_ctas) {
    // ---- producer ------------------------------------------------------------------------------------------------
    for (int j = blockIdx.x; j < blocks; j += push_ctas) {
      const int r0 = j * kTile, nrows = min(kTile, rows - r0);
      const size_t vecs = size_t(nrows) * K / 8;  // 16-byte vectors of this block (K % 8 == 0)
      const int4* src = reinterpret_cast<const int4*>(a + size_t(r0) * K);
      __nv_bfloat16* dst_local = my_ws + (size_t(rank) * rows + r0) * K;
      for (size_t i = threadIdx.x; i < vecs; i += kThreads) {
        const int4 v = ptx::ld_nc_v4(src + i);
        if (ctx.mc_base != nullptr) {
          ptx::multimem_st_v4(reinterpret_cast<int4*>(mc_ptr(ctx, dst_local)) + i, v);  // one store reaches all ranks
        } else {
          for (int p = 1; p < W; ++p) ptx::st_v4(reinterpret_cast<int4*>(symm_at(ctx, dst_local, (rank + p) % W)) + i, v);
        }
      }
      __syncthreads();  // every thread's stores are issued ...
      if (threadIdx.x == 0) {
        fence();  // ... and performed system-wide before the flag can be seen (release pattern)
        for (int p = 1; p < W; ++p) notify(ctx, flags + rank * flag_blocks + j, (rank + p) % W, epoch);
      }
    }
  } else {
    // ---- consumer ------------------------------------------------------------------------------------------------
    const int tiles_n = (N + kTile - 1) / kTile;
    const int tiles = W * blocks * tiles_n;
    const int consumers = gridDim.x - push_ctas;
    for (int t = blockIdx.x - push_ctas; t < tiles; t += consumers) {
      const int unit = t / tiles_n, n_blk = t % tiles_n;
      int shard, blk;
      if (unit < blocks) {  // own shard first
        shard = rank;
        blk = unit;
      } else {  // then remote blocks in arrival order
        const int u = unit - blocks;
        blk = u / (W - 1);
        shard = (rank + 1 + u % (W - 1)) % W;
      }
      const __nv_bfloat16* A = a;
      if (shard != rank) {
        if (threadIdx.x == 0) wait_ge_sys(flags + shard * flag_blocks + blk, epoch);  // acquire: the block has landed
        __syncthreads();
        A = my_ws + size_t(shard) * rows * K;
      }
      float acc[4][4];
      tile_gemm(A, K, rows, b, K, N, K, blk * kTile, n_blk * kTile, acc);
      const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
      for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
          const int r = blk * kTile + ty * 4 + i, col = n_blk * kTile + tx * 4 + j;
          if (r < rows && col < N) c[(size_t(shard) * rows + r) * N + col] = __float2bfloat16(acc[i][j]);
        }
    }
  }
  phase_end(ps, base, 2);  // the last CTA to leave advances the epoch for the next call
}

// ================================================================= tutorial 8: GEMM + ReduceScatter in ONE kernel
// Every rank holds a K-slice: partial = A[M, Kl] * B[N, Kl]^T; rank r must end up with out[M/W, N] = the sum over ranks
// of rows [r * M/W, (r+1) * M/W) of their partials.
//   phase 1: all ranks walk the tiles in the SAME owner-interleaved order (tile t belongs to owner t % W), store each
//            partial tile into their OWN symmetric buffer and flag the owner. Nothing but the flag crosses NVLink.
//   phase 2: the owner sums a tile as soon as all W flags of that tile are there: one multimem.ld_reduce per 16 bytes
//            makes the switch fetch the W copies and add them in fp32 (peer loads + adds without a multicast mapping).
// In this tutorial kernel phase 2 simply follows phase 1 in every CTA; in gemm_sm100.cuh two otherwise idle warps of
// every GEMM CTA run phase 2 WHILE the tensor core produces later tiles, and a third warp publishes the flags.
// part: [2][parity_elems] symmetric, parity_elems >= M * N and independent of this call's M (see the note on the workspace
// halves above); flags: [W][flag_blocks][N/64] symmetric with flag_blocks >= M/W/64 fixed for the buffer, zero initialised once.
__global__ void __launch_bounds__(kThreads) tut_gemm_rs_kernel(SymmCtx ctx, const __nv_bfloat16* a, const __nv_bfloat16* b,
                                                               __nv_bfloat16* out, __nv_bfloat16* part, size_t parity_elems,
                                                               uint32_t* flags, int flag_blocks, int M, int N, int Kl,
                                                               PhaseState ps) {
  const uint32_t base = phase_begin(ps);
  const uint32_t epoch = base + 2;
  const int parity = (epoch >> 1) & 1;
  const int W = ctx.world, rank = ctx.rank;
  const int rows = M / W;
  const int blocks = (rows + kTile - 1) / kTile, tiles_n = (N + kTile - 1) / kTile;
  __nv_bfloat16* my_part = part + size_t(parity) * parity_elems;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  // ---- phase 1: produce ------------------------------------------------------------------------------------------
  const int tiles = W * blocks * tiles_n;
  for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int owner = t % W, i = t / W;
    const int blk = i % blocks, n_blk = i / blocks;
    float acc[4][4];
    tile_gemm(a + size_t(owner) * rows * Kl, Kl, rows, b, Kl, N, Kl, blk * kTile, n_blk * kTile, acc);
    for (int ii = 0; ii < 4; ++ii)
      for (int j = 0; j < 4; ++j) {
        const int r = blk * kTile + ty * 4 + ii, col = n_blk * kTile + tx * 4 + j;
        if (r < rows && col < N) my_part[(size_t(owner) * rows + r) * N + col] = __float2bfloat16(acc[ii][j]);
      }
    __syncthreads();
    if (threadIdx.x == 0) {
      fence();  // my partial tile is visible system-wide before the owner can see the flag
      notify(ctx, flags + (size_t(rank) * flag_blocks + blk) * tiles_n + n_blk, owner, epoch);
    }
  }
  // ---- phase 2: reduce my segment ---------------------------------------------------------------------------------
  for (int u = blockIdx.x; u < blocks * tiles_n; u += gridDim.x) {
    const int blk = u % blocks, n_blk = u / blocks;
    if (threadIdx.x < W) wait_ge_sys(flags + (size_t(threadIdx.x) * flag_blocks + blk) * tiles_n + n_blk, epoch);
    __syncthreads();
    const int r0 = blk * kTile, c0 = n_blk * kTile;
    const int nrows = min(kTile, rows - r0), vpr = min(kTile, N - c0) / 8;  // N % 8 == 0
    for (int idx = threadIdx.x; idx < nrows * vpr; idx += kThreads) {
      const int r = idx / vpr, cv = idx % vpr;
      const size_t off = (size_t(rank) * rows + r0 + r) * N + c0 + cv * 8;
      int4 v;
      if (ctx.mc_base != nullptr) {
        v = ptx::multimem_ld_reduce_bf16x8(mc_ptr(ctx, my_part) + off);  // the switch adds the W copies
      } else {
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int p = 0; p < W; ++p) Pack<BF16>::add(s, ptx::ld_v4(symm_at(ctx, my_part, (rank + p) % W) + off));
        v = Pack<BF16>::cvt(s);
      }
      *reinterpret_cast<int4*>(out + (size_t(r0 + r)) * N + c0 + cv * 8) = v;
    }
  }
  phase_end(ps, base, 2);
}

}  // namespace tut
}  // namespace tdist

using namespace tdist;

extern "C" {
/* tutorials/07: C[W * rows, N] = all_gather(a[rows, K]) * b[N, K]^T. ws symmetric, ws_bytes >= 2 * W * rows * K * 2 (two
 * halves ws_bytes / 2 apart, whatever `rows` is), flags symmetric [W][flag_blocks] zeroed once (flag_blocks = ceil(largest rows
 * / 64), the same for every call on these buffers), phase/done device counters zeroed once. K % 8 == 0. */
int tdist_tutorial_ag_gemm(const tdist_symm_ctx* ctx, const void* a, const void* b, void* c, void* ws, size_t ws_bytes,
                           uint32_t* flags, int flag_blocks, uint32_t* phase, uint32_t* done, int rows, int N, int K,
                           int push_ctas, int ctas, void* stream) {
  if (K % 8 != 0 || push_ctas < 1 || ctas <= push_ctas || flag_blocks < (rows + tut::kTile - 1) / tut::kTile) {
    set_error("tutorial_ag_gemm: K % 8 == 0, 1 <= push_ctas < ctas, flag_blocks >= ceil(rows / 64)");
    return -1;
  }
  const SymmCtx cx = to_ctx(ctx);
  const size_t parity_elems = ws_bytes / 2 / sizeof(__nv_bfloat16) / 8 * 8;
  if (parity_elems < size_t(cx.world) * rows * K) {
    set_error("tutorial_ag_gemm: workspace too small (need 2 * W * rows * K elements)");
    return -1;
  }
  if (cx.world == 1) push_ctas = 0;  // nothing to push
  tut::tut_ag_gemm_kernel<<<ctas, tut::kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      cx, static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), static_cast<__nv_bfloat16*>(c),
      static_cast<__nv_bfloat16*>(ws), parity_elems, flags, flag_blocks, rows, N, K, push_ctas, PhaseState{phase, done});
  return check_launch("tutorial_ag_gemm");
}
/* tutorials/08: out[M/W, N] = reduce_scatter(a[M, Kl] * b[N, Kl]^T). part symmetric, part_bytes >= 2 * M * N * 2 (two
 * halves part_bytes / 2 apart, whatever M is), flags symmetric [W][flag_blocks][ceil(N/64)] zeroed once (flag_blocks =
 * ceil(largest M / W / 64), N the same for every call on these buffers). N % 8 == 0, M % W == 0. All CTAs must be co-resident (ctas <= SM count). */
int tdist_tutorial_gemm_rs(const tdist_symm_ctx* ctx, const void* a, const void* b, void* out, void* part, size_t part_bytes,
                           uint32_t* flags, int flag_blocks, uint32_t* phase, uint32_t* done, int M, int N, int Kl, int ctas,
                           void* stream) {
  const SymmCtx cx = to_ctx(ctx);
  if (N % 8 != 0 || M % cx.world != 0 || ctas < 1 || flag_blocks < (M / cx.world + tut::kTile - 1) / tut::kTile) {
    set_error("tutorial_gemm_rs: N % 8 == 0, M % world == 0, flag_blocks >= ceil(M / world / 64)");
    return -1;
  }
  const size_t parity_elems = part_bytes / 2 / sizeof(__nv_bfloat16) / 8 * 8;
  if (parity_elems < size_t(M) * N) {
    set_error("tutorial_gemm_rs: partial buffer too small (need 2 * M * N elements)");
    return -1;
  }
  tut::tut_gemm_rs_kernel<<<ctas, tut::kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      cx, static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), static_cast<__nv_bfloat16*>(out),
      static_cast<__nv_bfloat16*>(part), parity_elems, flags, flag_blocks, M, N, Kl, PhaseState{phase, done});
  return check_launch("tutorial_gemm_rs");
}
}
