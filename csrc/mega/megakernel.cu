// megakernel.cu — the persistent decode megakernel: one launch executes a whole decode step.
//
// Every CTA (one per SM) walks its static work queue; a task spins on the scoreboard slots of the ops it
// depends on, runs its body, and bumps its own op's slot. Scoreboard values are monotone (slot value after step
// s == s * tiles(op)), so nothing is reset between steps and the launch is CUDA-graph capturable.
// Reference dispatcher: mega_triton_kernel/core/code_generator.py:31-89; task bodies:
// mega_triton_kernel/kernels/{linear,mlp_fc1,norm,activation,elementwise,flash_attn,allreduce,barrier,prefetch}.py.
//
// B200 specifics: the decode linears are weight-streaming skinny GEMMs (M = batch <= 8 rows): the weights of all linear
// tasks of a CTA form ONE stream of TMA stages (3-D tiled loads, 128 B swizzle) into per-warp stage buffers that runs
// ahead across task boundaries — weights do not depend on activations, so the HBM latency of every task start hides
// behind the scoreboard wait — and a stage is consumed with ldmatrix + mma.sync.m16n8k16 (weights as the M operand, the
// batch as N; see LINEAR below). RMSNorm is folded into the consumer linear's prologue, SiLU*up into fc1's epilogue and
// the residual add into the all-reduce, which removes three scoreboard hops per layer. The all-reduce is NVLS
// multimem.ld_reduce with a per-tile cross-rank flag exchange.
// The whole kernel also runs on the CPU under ThreadSanitizer with simulated ranks (csrc/sim/mega_scenarios.h).
#include <cuda_bf16.h>
#include <math_constants.h>

#include "../kernels/common.cuh"
#include "mega.h"

namespace tdist {
namespace mega {

using bf16 = __nv_bfloat16;

__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = ptx::unpack_bf16x2(u[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ int4 pack8(const float (&f)[8]) {
  int4 r;
  r.x = ptx::pack_bf16x2(f[0], f[1]);
  r.y = ptx::pack_bf16x2(f[2], f[3]);
  r.z = ptx::pack_bf16x2(f[4], f[5]);
  r.w = ptx::pack_bf16x2(f[6], f[7]);
  return r;
}

__device__ __forceinline__ float block_sum(float v, float* scratch) {
  v = ptx::warp_reduce_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) scratch[w] = v;
  __syncthreads();
  float r = l < (kThreads >> 5) ? scratch[l] : 0.f;
  r = ptx::warp_reduce_sum(r);
  return r;
}

__device__ __forceinline__ void wait_slot(const uint32_t* sb, int slot, uint32_t target) {
  if (slot < 0) return;
  // (polling with ld.volatile + one trailing fence instead of ld.acquire was measured slower: 4.60 vs 4.27 ms/token)
  TDIST_SPIN_GUARD_DECL;
  uint32_t v;
  while (static_cast<int32_t>((v = ptx::ld_acquire_gpu(sb + slot)) - target) < 0) {
    TDIST_SPIN_GUARD(sb + slot, v, target)
  }
}

// ------------------------------------------------------------------ LINEAR (skinny GEMM, M <= 8 rows)
// i0 bit0: silu pairing (W = [2I, K], out = silu(gate) * up, N = I); i1: ldy; i2: out parity stride (bytes, 0 = none)
//
// Decode linears are weight-streaming: every weight byte is used once, so the job is to keep as many bytes in flight
// from HBM as shared memory can hold (loaded HBM latency on B200 is several microseconds: full bandwidth needs
// ~200 KB per SM in flight) with as few instructions and copy requests per byte as possible.
//
// The weights of a tile (cols_per_tile consecutive rows of W[N, K]) are cut into row blocks of R <= 24 rows and every
// block into STAGES of R rows x (256 .. 768) k-elements (~12 KB whatever R is). One stage is ONE TMA request
// (cp.async.bulk.tensor.3d on the view [k % 64][row][k / 64] of W, box {64, R, nsub}, 128 B swizzle so that
// ldmatrix is bank-conflict free) completed on one mbarrier. The stages of ALL linear tasks of this CTA form one sequence numbered from 0 (static table built by the
// host, mega.h StreamEntry, copied to shared memory at kernel start).
//
// Shared memory holds `slots` stage buffers and warp w (< slots) OWNS buffer w: it consumes stage s = w (mod slots)
// and, as soon as its ldmatrix reads are done, issues stage s + slots into the same buffer itself. There is no
// producer warp, no empty barrier and no coupling between warps inside a row block; weights never depend on
// activations, so the stream runs ahead across task boundaries on its own: the first stages of the NEXT linear task
// are in flight while the current task's epilogue runs, while small tasks (attention, norms) execute and while the
// scoreboard wait spins.
//
// A stage is consumed with mma.sync.m16n8k16 (bf16 x bf16 -> fp32): the weight rows are the M dimension of the MMA,
// the batch (<= 8 activation rows) is its N dimension, so one ldmatrix + one MMA retire 512 B of weights and a batch
// of 8 costs the same as a batch of 1. Each warp keeps fp32 accumulator tiles for the current row block; the warps'
// partials are summed through shared memory once per block (one __syncthreads per block).
// (Measured on the way here: per-row 1-D bulk copies are limited by the TMA request rate, an unswizzled {256, R} box
// by 8-way conflicted ldmatrix, and a ring shared by all warps by the per-stage handshake latency.)
struct Ring {
  uint8_t* buf;      // slots * kSlotBytes
  uint64_t* full;    // [slots] : 1 arrival (+ tx bytes)
  const StreamEntry* tab;
  uint32_t n_ent;
  uint32_t slots;
  // per-warp stream cursor (warp-uniform registers)
  uint32_t issued;   // id of the stage in flight / next to be consumed by this warp (0xffffffff: none)
  uint32_t e_idx, e_first, e_total;  // table entry containing `issued`
  uint32_t uses;     // completed uses of this warp's buffer (mbarrier parity)
  // second cursor for the L2 prefetch of this warp's stages further ahead (see prefetch_stage)
  uint32_t pf_next, pf_idx, pf_first, pf_total;
  // optional per-warp cycle accounting of the linear main loop (KernelParams::trace != nullptr):
  // [0] waiting for a stage to land, [1] ldmatrix + MMA, [2] issuing the next stage, [3] block-end barrier + sum
  bool prof;
  uint32_t cyc[4];
};
__device__ __forceinline__ uint32_t clk() { return uint32_t(clock64()); }

__device__ __forceinline__ void stage_coords(const StreamEntry& e, uint32_t stage, int& c0, int& c1) {
  uint32_t q = stage - e.first;
  const uint32_t per_blob = uint32_t(e.nblk) * e.kboxes;
  const bool second = q >= per_blob;
  if (second) q -= per_blob;
  // q / kboxes without an integer division (q < 2^16, kboxes <= 400: the float quotient of (q + 0.5) is exact enough)
  const uint32_t blk = uint32_t((float(q) + 0.5f) * __frcp_rn(float(e.kboxes)));
  const uint32_t kb = q - blk * e.kboxes;
  c0 = int(kb * uint32_t(e.kstage));
  c1 = int(e.row0 + (second ? e.pair_rows : 0u) + blk * e.R);
}

// L2 prefetch of this warp's future stages up to (excluding) `limit` (warp-collective). The stage buffers alone hold
// ~190 KB per SM; at TP8 a linear task mostly waits (dependencies, activation staging) and then has to pull the rest
// of its tile through that window at HBM latency. Two uses: one buffer-round ahead of every stage load (keeps the
// steady-state loads L2 hits without flooding L2), and up to kIdlePrefetchRounds rounds ahead while a CTA waits on
// the scoreboard (HBM is idle then, the whole next tile fits in L2).
__device__ __forceinline__ void prefetch_upto(Ring& r, uint32_t limit) {
  while (r.pf_next < limit) {
    while (r.pf_idx < r.n_ent && r.pf_next >= r.pf_first + r.pf_total) {
      ++r.pf_idx;
      if (r.pf_idx < r.n_ent) {
        r.pf_first = r.tab[r.pf_idx].first;
        r.pf_total = r.tab[r.pf_idx].total;
      }
    }
    if (r.pf_idx >= r.n_ent) {
      r.pf_next = 0xfffffff0u;
      return;
    }
    if ((threadIdx.x & 31) == 0) {
      const StreamEntry e = r.tab[r.pf_idx];
      int c0, c1;
      stage_coords(e, r.pf_next, c0, c1);
      ptx::tma_prefetch_3d(e.tmap, 0, c1, c0 >> 6);
    }
    r.pf_next += r.slots;
  }
}

// Issue stage `stage` into this warp's buffer (warp-collective; lane 0 talks to the TMA). The caller guarantees that
// the warp's previous reads of the buffer are complete.
__device__ __forceinline__ void issue_stage(Ring& r, uint32_t stage) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  while (r.e_idx < r.n_ent && stage >= r.e_first + r.e_total) {
    ++r.e_idx;
    if (r.e_idx < r.n_ent) {
      r.e_first = r.tab[r.e_idx].first;
      r.e_total = r.tab[r.e_idx].total;
    }
  }
  if (r.e_idx >= r.n_ent) {
    r.issued = 0xffffffffu;  // the stream is finished for this warp
    return;
  }
  r.issued = stage;
  if (lane == 0) {
    const StreamEntry e = r.tab[r.e_idx];
    int c0, c1;
    stage_coords(e, stage, c0, c1);
    ptx::fence_proxy_async_shared();  // order the generic-proxy reads of the old contents before the async-proxy writes
    // ONE request: box {64 k, R rows, nsub k-slabs} of the 3-D view [k % 64][row][k / 64]; lands as [slab][row][128 B]
    const int nsub = e.kstage >> 6;
    ptx::mbar_arrive_expect_tx(&r.full[warp], uint32_t(e.R) * 128u * uint32_t(nsub));
    ptx::tma_load_3d(r.buf + size_t(warp) * kSlotBytes, e.tmap, &r.full[warp], 0, c1, c0 >> 6);
  }
  if (kPrefetchRounds > 0) prefetch_upto(r, stage + kPrefetchRounds * r.slots + 1);
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&a)[4], const void* smem_ptr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3])
               : "r"(ptx::smem_u32(smem_ptr)));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void linear_body(const OpArgs& a, int tile, uint8_t* smem, uint32_t step, const uint32_t* sb,
                                            const Task& t, uint32_t dep_target0, uint32_t dep_target1, Ring& r,
                                            uint32_t& stage_base, uint32_t lin_k, uint64_t* tr) {
  constexpr int kWarps = kThreads / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int K = a.K, kv = K >> 3, M = a.M, cpt = a.cols_per_tile;
  const bool pair = (a.i0 & 1) != 0;
  const int c_beg = tile * cpt, c_end = min(c_beg + cpt, a.N);
  const int ncols = c_end - c_beg;
  const LinearGeom ge = linear_geom(cpt, K);                                  // same split as the host (StreamEntry)
  const int xpitch = ge.kboxes * ge.kstage + 8;                               // elements: zero padded to whole stages, +16 B
  bf16* xs = reinterpret_cast<bf16*>(smem);                                  // [M][xpitch] (normalised) activations
  const bf16* nw = static_cast<const bf16*>(a.in2);
  bf16* nws = xs + size_t(M) * xpitch;                                        // [K] norm weight (only when fused)
  float* acc = reinterpret_cast<float*>(nws + (nw != nullptr ? K : 0));       // [2][nblk][2][16][M] block sums
  const int tile_vals = 32 * M;                                               // (2 row groups x 16 rows) x M per block
  float* red = acc + 2 * size_t(ge.nblk) * tile_vals;                         // [2][kWarps][2][16][M] partials
  float* scratch = red + 2 * kWarps * tile_vals;                              // 32 floats for block_sum

  // (1) everything that does not depend on the producer op is done before the scoreboard wait: norm weight into
  // shared memory, zero padding of x. (This task's first stages are already in flight, see Ring.)
  if (nw != nullptr)
    for (int i = threadIdx.x; i < kv; i += kThreads)
      *reinterpret_cast<int4*>(nws + size_t(i) * 8) = ptx::ld_nc_v4(nw + size_t(i) * 8);
  for (int m = 0; m < M; ++m)
    for (int i = K + threadIdx.x; i < xpitch; i += kThreads) xs[size_t(m) * xpitch + i] = __float2bfloat16(0.f);
  // HBM is idle while this CTA waits: pull this warp's next stages (this tile and at most the next linear tile) into L2
  if (kIdlePrefetchRounds > 0 && uint32_t(warp) < r.slots && r.issued != 0xffffffffu) {
    const uint32_t this_total = (pair ? 2u : 1u) * uint32_t(ge.nblk) * uint32_t(ge.kboxes);
    const uint32_t next_total = lin_k + 1 < r.n_ent ? r.tab[lin_k + 1].total : 0u;
    prefetch_upto(r, min(r.issued + kIdlePrefetchRounds * r.slots + 1, stage_base + this_total + next_total));
  }
  // (2) wait for the activations
  if (threadIdx.x == 0) {
    wait_slot(sb, t.dep0, dep_target0);
    wait_slot(sb, t.dep1, dep_target1);
    if (tr != nullptr) tr[2] = ptx::globaltimer();  // dependencies satisfied
  }
  __syncthreads();
  // (3) stage x (optionally RMS-normalised) in shared memory
  const bf16* x = static_cast<const bf16*>(a.in0);
  const float eps = a.f0;
  for (int m = 0; m < M; ++m) {
    float ss = 0.f;
    for (int i = threadIdx.x; i < kv; i += kThreads) {
      const int4 raw = ptx::ld_v4(x + size_t(m) * K + size_t(i) * 8);
      *reinterpret_cast<int4*>(xs + size_t(m) * xpitch + size_t(i) * 8) = raw;
      if (nw != nullptr) {
        float f[8];
        unpack8(raw, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) ss = fmaf(f[j], f[j], ss);
      }
    }
    if (nw != nullptr) {
      const float tot = block_sum(ss, scratch);
      const float inv = rsqrtf(tot / K + eps);
      for (int i = threadIdx.x; i < kv; i += kThreads) {
        float f[8], w[8];
        unpack8(*reinterpret_cast<const int4*>(xs + size_t(m) * xpitch + size_t(i) * 8), f);
        unpack8(*reinterpret_cast<const int4*>(nws + size_t(i) * 8), w);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = __bfloat162float(__float2bfloat16(f[j] * inv)) * w[j];  // round like the unfused norm
        *reinterpret_cast<int4*>(xs + size_t(m) * xpitch + size_t(i) * 8) = pack8(f);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && tr != nullptr) tr[3] = ptx::globaltimer();  // activations staged

  // (4) consume the weight stream with tensor-core MMAs
  // A fragment addressing (ldmatrix.x4): lane l supplies row (l & 7) + 8 * ((l >> 3) & 1), k offset 8 * (l >> 4).
  // Buffer layout: [nsub slabs][R rows][128 B], the 16-byte chunks of every 128 B line XOR-swizzled with (line & 7)
  // by the TMA (line = slab * R + row; the buffer is 1024 B aligned).
  const uint32_t a_row = uint32_t((lane & 7) + ((lane >> 3) & 1) * 8);  // row inside a 16-row group
  const uint32_t a_hi = uint32_t(lane >> 4);                            // which 16-byte half of the k-slice
  const uint8_t* sbuf = r.buf + size_t(warp) * kSlotBytes;
  // B fragment: this lane holds x[n = lane / 4][k0 + 2 (lane % 4) + {0, 1}] and the same at k0 + 8
  const int bn = lane >> 2;
  const bf16* xb = xs + size_t(bn < M ? bn : 0) * xpitch + (lane & 3) * 2;
  const uint32_t bmask = bn < M ? 0xffffffffu : 0u;  // batch rows beyond M contribute zeros
  const bool two_groups = ge.R > 16;
  const bool consumer = uint32_t(warp) < r.slots;
  int rbuf = 0;
  uint32_t blk_first = stage_base;  // id of the first stage of the current (blob, row block)
  for (int blob = 0; blob < (pair ? 2 : 1); ++blob) {
    for (int blk = 0; blk < ge.nblk; ++blk, blk_first += ge.kboxes) {
      // two accumulator tiles per row group (even / odd k-slices): halves the dependent-MMA chain
      float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
      float d0b[4] = {0.f, 0.f, 0.f, 0.f}, d1b[4] = {0.f, 0.f, 0.f, 0.f};
      if (consumer) {
        // this warp's stages of the block: ids = warp (mod slots)
        uint32_t kb = (uint32_t(warp) + r.slots - blk_first % r.slots) % r.slots;
        for (; kb < uint32_t(ge.kboxes); kb += r.slots) {
          // stage blk_first + kb is the one in flight in this warp's buffer (issued after the previous one was consumed)
          uint32_t c0 = r.prof ? clk() : 0u;
          ptx::mbar_wait(&r.full[warp], r.uses & 1u);
          ++r.uses;
          uint32_t c1 = r.prof ? clk() : 0u;
          // slabs of 64 k-elements that hold real data (the activations are zero padded to whole stages, so a
          // partly filled slab needs no special case; K is a multiple of 64)
          const int nslab = min(ge.nsub, (K - int(kb) * ge.kstage) >> 6);
          const bf16* xk = xb + int(kb) * ge.kstage;
          uint32_t line = a_row;  // 128 B line of this lane's row in slab 0; slab j starts at line j * R
#pragma unroll 2
          for (int j = 0; j < nslab; ++j, line += ge.R, xk += 64) {
            const uint8_t* lp = sbuf + line * 128u;
            const uint32_t ph = line & 7u;  // swizzle phase of the line (rows + 16 have the same one)
#pragma unroll
            for (int s4 = 0; s4 < 4; ++s4) {
              const uint32_t b0 = *reinterpret_cast<const uint32_t*>(xk + s4 * 16) & bmask;
              const uint32_t b1 = *reinterpret_cast<const uint32_t*>(xk + s4 * 16 + 8) & bmask;
              const uint8_t* ap = lp + ((uint32_t(s4 * 2) + a_hi) ^ ph) * 16u;
              uint32_t af[4];
              ldmatrix_x4(af, ap);
              mma_bf16_16816(s4 & 1 ? d0b : d0, af, b0, b1);
              if (two_groups) {
                ldmatrix_x4(af, ap + 16 * 128);
                mma_bf16_16816(s4 & 1 ? d1b : d1, af, b0, b1);
              }
            }
          }
          __syncwarp();
          uint32_t c2 = r.prof ? clk() : 0u;
          issue_stage(r, blk_first + kb + r.slots);  // refill this buffer right away (possibly for a later task)
          if (r.prof) {
            r.cyc[0] += c1 - c0;
            r.cyc[1] += c2 - c1;
            r.cyc[2] += clk() - c2;
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        d0[i] += d0b[i];
        d1[i] += d1b[i];
      }
      const uint32_t c3 = r.prof ? clk() : 0u;
      // block done: sum the warps' partial tiles. d[0..1] = (row lane/4, n 2*(lane%4)+{0,1}), d[2..3] = row + 8.
      {
        float* rb = red + (size_t(rbuf) * kWarps + warp) * tile_vals;
        const int n0 = (lane & 3) * 2, row = lane >> 2;
        if (n0 < M) {
          rb[row * M + n0] = d0[0];
          rb[(row + 8) * M + n0] = d0[2];
          rb[(16 + row) * M + n0] = d1[0];
          rb[(24 + row) * M + n0] = d1[2];
        }
        if (n0 + 1 < M) {
          rb[row * M + n0 + 1] = d0[1];
          rb[(row + 8) * M + n0 + 1] = d0[3];
          rb[(16 + row) * M + n0 + 1] = d1[1];
          rb[(24 + row) * M + n0 + 1] = d1[3];
        }
      }
      __syncthreads();
      if (int(threadIdx.x) < tile_vals) {
        const float* rs = red + size_t(rbuf) * kWarps * tile_vals + threadIdx.x;
        float sum = 0.f;
        for (uint32_t w = 0; w < r.slots; ++w) sum += rs[w * tile_vals];
        acc[(size_t(blob) * ge.nblk + blk) * tile_vals + threadIdx.x] = sum;  // [(blob, block)][row 0..31][m]
      }
      rbuf ^= 1;  // the next block's partials go to the other buffer; its barrier orders the reuse of this one
      if (r.prof) r.cyc[3] += clk() - c3;
    }
  }
  stage_base = blk_first;
  __syncthreads();
  if (threadIdx.x == 0 && tr != nullptr) tr[4] = ptx::globaltimer();  // weight stream consumed

  // (5) epilogue: activation pairing, bias, store
  char* out_base = static_cast<char*>(a.out0) + (a.i2 ? size_t(step & 1) * a.i2 : 0);
  bf16* y = reinterpret_cast<bf16*>(out_base);
  const bf16* bias = static_cast<const bf16*>(a.in3);
  const int ldy = a.i1;
  for (int idx = threadIdx.x; idx < ncols * M; idx += kThreads) {
    const int m = idx / ncols, c = idx - m * ncols;
    const int blk = c / ge.R, rr = c - blk * ge.R;
    const size_t off = size_t(blk) * tile_vals + rr * M + m;
    float v = acc[off];
    if (pair) {
      const float u = acc[size_t(ge.nblk) * tile_vals + off];
      // round gate/up to bf16 first: matches the unfused fc1 -> silu_mul sequence bit for bit in the common case
      v = __bfloat162float(__float2bfloat16(v));
      const float uu = __bfloat162float(__float2bfloat16(u));
      v = v / (1.f + __expf(-v)) * uu;
    }
    if (bias != nullptr) v += __bfloat162float(bias[c_beg + c]);
    y[size_t(m) * ldy + c_beg + c] = __float2bfloat16(v);
  }
}

// ------------------------------------------------------------------ RMS_NORM (tile = row), optional residual add
__device__ __forceinline__ void rmsnorm_body(const OpArgs& a, int tile, uint8_t* smem) {
  const int K = a.K, kv = K >> 3;
  float* row = reinterpret_cast<float*>(smem);
  float* scratch = row + K;
  const bf16* x = static_cast<const bf16*>(a.in0) + size_t(tile) * K;
  bf16* res = a.in3 ? static_cast<bf16*>(const_cast<void*>(a.in3)) + size_t(tile) * K : nullptr;
  float ss = 0.f;
  for (int i = threadIdx.x; i < kv; i += kThreads) {
    float f[8];
    unpack8(ptx::ld_v4(x + size_t(i) * 8), f);
    if (res) {
      float r[8];
      unpack8(ptx::ld_v4(res + size_t(i) * 8), r);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += r[j];
      const int4 p = pack8(f);
      *reinterpret_cast<int4*>(res + size_t(i) * 8) = p;
      unpack8(p, f);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      row[i * 8 + j] = f[j];
      ss = fmaf(f[j], f[j], ss);
    }
  }
  const float inv = rsqrtf(block_sum(ss, scratch) / K + a.f0);
  const bf16* w = static_cast<const bf16*>(a.in2);
  bf16* y = static_cast<bf16*>(a.out0) + size_t(tile) * K;
  for (int i = threadIdx.x; i < kv; i += kThreads) {
    float wv[8], o[8];
    unpack8(ptx::ld_nc_v4(w + size_t(i) * 8), wv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = __bfloat162float(__float2bfloat16(row[i * 8 + j] * inv)) * wv[j];
    *reinterpret_cast<int4*>(y + size_t(i) * 8) = pack8(o);
  }
}

// ------------------------------------------------------------------ elementwise
__device__ __forceinline__ void silu_mul_body(const OpArgs& a, int tile) {
  const int inter = a.N, iv = inter >> 3;
  const long long total = (long long)a.M * iv;
  const bf16* gu = static_cast<const bf16*>(a.in0);
  bf16* out = static_cast<bf16*>(a.out0);
  for (long long idx = (long long)tile * kThreads + threadIdx.x; idx < total; idx += (long long)a.tiles * kThreads) {
    const long long r = idx / iv;
    const int c = int(idx % iv);
    float g[8], u[8], o[8];
    unpack8(ptx::ld_v4(gu + r * 2 * inter + size_t(c) * 8), g);
    unpack8(ptx::ld_v4(gu + r * 2 * inter + inter + size_t(c) * 8), u);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = g[j] / (1.f + __expf(-g[j])) * u[j];
    *reinterpret_cast<int4*>(out + r * inter + size_t(c) * 8) = pack8(o);
  }
}
__device__ __forceinline__ void add_body(const OpArgs& a, int tile) {
  const long long nv = (long long)a.N >> 3;
  const int4* x = static_cast<const int4*>(a.in0);
  const int4* y = static_cast<const int4*>(a.in1);
  int4* o = static_cast<int4*>(a.out0);
  for (long long i = (long long)tile * kThreads + threadIdx.x; i < nv; i += (long long)a.tiles * kThreads) {
    float p[8], q[8];
    unpack8(ptx::ld_v4(x + i), p);
    unpack8(ptx::ld_v4(y + i), q);
#pragma unroll
    for (int j = 0; j < 8; ++j) p[j] += q[j];
    o[i] = pack8(p);
  }
}

// ------------------------------------------------------------------ q/k norm + RoPE + KV append (head_dim 128)
// in0 qkv [B, (Hq+2Hkv)*128]; in1/in2 k/v cache; in3/in4 q/k norm weights; in5 cos_sin fp32 [max_pos,128];
// out0 q_out [B,Hq,128]; out1 positions (int32 [B]); out2 block_table (int32 [B,max_pages]) or null.
// i0 Hq, i1 Hkv, i2 page_size, i3 max_pages; l0/l1/l2 batch(page)/seq/head strides of the cache.
__device__ __forceinline__ void qk_rope_body(const OpArgs& a, int tile) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Hq = a.i0, Hkv = a.i1, heads = Hq + 2 * Hkv;
  const int total = a.M * heads;
  const int* positions = static_cast<const int*>(a.out1);
  const int* block_table = static_cast<const int*>(a.out2);
  for (int item = tile * (kThreads >> 5) + warp; item < total; item += a.tiles * (kThreads >> 5)) {
    const int b = item / heads, h = item % heads;
    const bf16* src = static_cast<const bf16*>(a.in0) + (size_t(b) * heads + h) * 128 + lane * 4;
    const uint2 raw = *reinterpret_cast<const uint2*>(src);
    float f[4];
    {
      const float2 p = ptx::unpack_bf16x2(raw.x), q = ptx::unpack_bf16x2(raw.y);
      f[0] = p.x, f[1] = p.y, f[2] = q.x, f[3] = q.y;
    }
    const bool is_q = h < Hq, is_k = !is_q && h < Hq + Hkv;
    const int pos = positions[b];
    if (is_q || is_k) {
      const bf16* nw = static_cast<const bf16*>(is_q ? a.in3 : a.in4);
      if (nw != nullptr) {
        float ss = f[0] * f[0] + f[1] * f[1] + f[2] * f[2] + f[3] * f[3];
        ss = ptx::warp_reduce_sum(ss);
        const float inv = rsqrtf(ss / 128.f + a.f0);
        const uint2 wr = *reinterpret_cast<const uint2*>(nw + lane * 4);
        const float2 wa = ptx::unpack_bf16x2(wr.x), wb = ptx::unpack_bf16x2(wr.y);
        f[0] = __bfloat162float(__float2bfloat16(f[0] * inv)) * wa.x;
        f[1] = __bfloat162float(__float2bfloat16(f[1] * inv)) * wa.y;
        f[2] = __bfloat162float(__float2bfloat16(f[2] * inv)) * wb.x;
        f[3] = __bfloat162float(__float2bfloat16(f[3] * inv)) * wb.y;
#pragma unroll
        for (int i = 0; i < 4; ++i) f[i] = __bfloat162float(__float2bfloat16(f[i]));
      }
      const float* cs = static_cast<const float*>(a.in5) + size_t(pos) * 128 + (lane & 15) * 4;
      const float4 c = *reinterpret_cast<const float4*>(cs), s = *reinterpret_cast<const float4*>(cs + 64);
      const float cc[4] = {c.x, c.y, c.z, c.w}, sn[4] = {s.x, s.y, s.z, s.w};
      float o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float other = __shfl_xor_sync(0xffffffffu, f[i], 16);
        o[i] = lane < 16 ? f[i] * cc[i] - other * sn[i] : f[i] * cc[i] + other * sn[i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) f[i] = o[i];
    }
    uint2 packed;
    packed.x = ptx::pack_bf16x2(f[0], f[1]);
    packed.y = ptx::pack_bf16x2(f[2], f[3]);
    if (is_q) {
      *reinterpret_cast<uint2*>(static_cast<bf16*>(a.out0) + (size_t(b) * Hq + h) * 128 + lane * 4) = packed;
    } else {
      const int kvh = is_k ? h - Hq : h - Hq - Hkv;
      size_t off;
      if (block_table != nullptr) {
        const int page = block_table[size_t(b) * a.i3 + pos / a.i2];
        off = size_t(page) * a.l0 + size_t(pos % a.i2) * a.l1 + size_t(kvh) * a.l2;
      } else {
        off = size_t(b) * a.l0 + size_t(pos) * a.l1 + size_t(kvh) * a.l2;
      }
      bf16* dst = static_cast<bf16*>(const_cast<void*>(is_k ? a.in1 : a.in2)) + off + lane * 4;
      *reinterpret_cast<uint2*>(dst) = packed;
    }
  }
}

// ------------------------------------------------------------------ attention: split-KV decode (head_dim 128)
// in0 q [B,Hq,128]; in1/in2 caches; in3 block_table; in4 positions (kv_len = pos + 1); out0 partial [B,Hq,S,129].
// i0 Hq, i1 Hkv, i2 page_size, i3 max_pages, i4 splits; f0 scale. tile = (b * Hkv + kvh) * S + split.
// FUSED (op ATTN_FUSED): the task also does what the reference runs as separate ops before and after it —
//  * q-norm + RoPE of its G query heads straight from the qkv projection (each split recomputes them: 8 x 128 values),
//    k-norm + RoPE + cache write of the NEW token by the split that owns the last position;
//  * the LSE combine over the splits, by whichever split of the (batch, kv head) finishes last (a monotone arrival
//    counter per (batch, kv head) behind the partial buffer; its value after step s is s * S).
// That removes two scoreboard hops and two small tasks from every layer of the decode step.
// FUSED operands: in0 qkv [B,(Hq+2Hkv)*128]; in5 cos_sin fp32 [max_pos,128]; out1 final out [B,Hq,128] bf16;
// out2 q/k norm weights bf16 [2][128] or null; f1 norm eps; counters = uint32 [B*Hkv] behind out0's B*Hq*S*129 floats.
// The number of query heads per pass (G <= 8) is a run-time value: one instantiation per FUSED instead of four keeps
// the interpreter's code small (with eight instantiations the kernel grew from 190 KB to 360 KB of SASS and every
// other task type, including the linear main loop, got measurably slower).
template <bool FUSED>
__device__ __forceinline__ void attn_split_body(const OpArgs& a, int tile, uint8_t* smem, int G, int g_total, int g_off,
                                                const uint32_t* wait_sb, const Task& wt, uint32_t tgt0, uint32_t tgt1,
                                                uint64_t* tr, uint32_t step) {
  // `wait_sb` != nullptr: this call also performs the task's scoreboard wait (first pass over the head group).
  // One warp per key position (lane = 4 of the 128 head dims), all G query heads of the KV group at once: the
  // per-thread state is G x 4 output accumulators + G (max, sum) pairs, which fits in registers for G = 8 (the first
  // version used half-warps with 8 dims per lane and spilled). The scaled queries sit in shared memory.
  constexpr int D = 128, NW = kThreads / 32, GM = 8;
  const int S = a.i4, Hq = a.i0, Hkv = a.i1;
  const int split = tile % S, kvh = (tile / S) % Hkv, b = tile / (S * Hkv);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int kv_len = static_cast<const int*>(a.in4)[b] + 1;
  const int per_split = (kv_len + S - 1) / S;
  const int s_beg = split * per_split, s_end = min(s_beg + per_split, kv_len);
  const int h0 = kvh * g_total + g_off;
  const bf16* q = static_cast<const bf16*>(a.in0);
  const bf16* kc = static_cast<const bf16*>(a.in1);
  const bf16* vc = static_cast<const bf16*>(a.in2);
  const int* block_table = static_cast<const int*>(a.in3);
  float* sh_m = reinterpret_cast<float*>(smem);  // [NW][G]
  float* sh_l = sh_m + NW * G;
  float* sh_o = sh_l + NW * G;                   // [NW][G][D]
  float* qs = sh_o + NW * G * D;                 // [G][D] queries, pre-scaled by sm_scale * log2(e)
  auto kv_off = [&](int pos) -> size_t {
    if (block_table != nullptr) {
      const int page = block_table[size_t(b) * a.i3 + pos / a.i2];
      return size_t(page) * a.l0 + size_t(pos % a.i2) * a.l1 + size_t(kvh) * a.l2;
    }
    return size_t(b) * a.l0 + size_t(pos) * a.l1 + size_t(kvh) * a.l2;
  };
  constexpr int UN = 2;
  // K/V of every position but the newest one were written by earlier steps: this warp's first rows are fetched BEFORE
  // the dependency wait (the block-table -> K/V chain is two dependent HBM round trips otherwise on the critical path)
  uint2 kr[UN], vr[UN];
  bool pre[UN];
  const int p_first = s_beg + warp * UN;
#pragma unroll
  for (int u = 0; u < UN; ++u) {
    pre[u] = wait_sb != nullptr && p_first + u < s_end && p_first + u < kv_len - 1;
    if (pre[u]) {
      const size_t off = kv_off(p_first + u) + lane * 4;
      kr[u] = *reinterpret_cast<const uint2*>(kc + off);
      vr[u] = *reinterpret_cast<const uint2*>(vc + off);
    }
  }
  // FUSED: everything the q / k preparation needs except the projection itself is known before the producer op
  // finishes (position, cos/sin row, norm weights): fetch it now, these loads are DRAM misses
  const int f_heads = Hq + 2 * Hkv, f_pos = kv_len - 1;
  const bool f_owner = s_beg <= f_pos && f_pos < s_end;
  const int f_role = !FUSED ? 3 : warp < G ? 0 : (f_owner && warp == G) ? 1 : (f_owner && warp == G + 1) ? 2 : 3;
  float4 f_cos = make_float4(0.f, 0.f, 0.f, 0.f), f_sin = f_cos;
  uint2 f_nw = make_uint2(0u, 0u);
  const bool f_has_nw = FUSED && a.out2 != nullptr;
  if (FUSED && f_role < 2) {
    const float* cs = static_cast<const float*>(a.in5) + size_t(f_pos) * D + (lane & 15) * 4;
    f_cos = *reinterpret_cast<const float4*>(cs);
    f_sin = *reinterpret_cast<const float4*>(cs + 64);
    if (f_has_nw) f_nw = *reinterpret_cast<const uint2*>(static_cast<const bf16*>(a.out2) + f_role * D + lane * 4);
  }
  if (wait_sb != nullptr && threadIdx.x == 0) {
    wait_slot(wait_sb, wt.dep0, tgt0);
    wait_slot(wait_sb, wt.dep1, tgt1);
    if (tr != nullptr) tr[2] = ptx::globaltimer();  // dependencies satisfied
  }
  __syncthreads();
  if (!FUSED) {
    for (int i = threadIdx.x; i < G * D / 8; i += kThreads) {
      const int g = i / (D / 8), c = i % (D / 8);
      float f[8];
      unpack8(ptx::ld_v4(q + (size_t(b) * Hq + h0 + g) * D + c * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) qs[g * D + c * 8 + j] = f[j] * (a.f0 * 1.4426950408889634f);
    }
  } else {
    // warp g < G: query head h0 + g; warps G / G + 1 of the split that owns the newest position: its K / V head
    const int heads = f_heads, pos = f_pos, role = f_role;
    if (role < 3) {
      const int h = role == 0 ? h0 + warp : role == 1 ? Hq + kvh : Hq + Hkv + kvh;
      const uint2 raw = *reinterpret_cast<const uint2*>(static_cast<const bf16*>(a.in0) + (size_t(b) * heads + h) * D + lane * 4);
      float f[4];
      {
        const float2 p0 = ptx::unpack_bf16x2(raw.x), p1 = ptx::unpack_bf16x2(raw.y);
        f[0] = p0.x, f[1] = p0.y, f[2] = p1.x, f[3] = p1.y;
      }
      if (role < 2) {
        if (f_has_nw) {
          float ss = f[0] * f[0] + f[1] * f[1] + f[2] * f[2] + f[3] * f[3];
          ss = ptx::warp_reduce_sum(ss);
          const float inv = rsqrtf(ss / 128.f + a.f1);
          const float2 wa = ptx::unpack_bf16x2(f_nw.x), wb = ptx::unpack_bf16x2(f_nw.y);
          f[0] = __bfloat162float(__float2bfloat16(f[0] * inv)) * wa.x;
          f[1] = __bfloat162float(__float2bfloat16(f[1] * inv)) * wa.y;
          f[2] = __bfloat162float(__float2bfloat16(f[2] * inv)) * wb.x;
          f[3] = __bfloat162float(__float2bfloat16(f[3] * inv)) * wb.y;
#pragma unroll
          for (int i = 0; i < 4; ++i) f[i] = __bfloat162float(__float2bfloat16(f[i]));
        }
        const float cc[4] = {f_cos.x, f_cos.y, f_cos.z, f_cos.w}, sn[4] = {f_sin.x, f_sin.y, f_sin.z, f_sin.w};
        float r[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float other = __shfl_xor_sync(0xffffffffu, f[i], 16);
          r[i] = lane < 16 ? f[i] * cc[i] - other * sn[i] : f[i] * cc[i] + other * sn[i];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) f[i] = __bfloat162float(__float2bfloat16(r[i]));  // the unfused op stores bf16
      }
      if (role == 0) {
        const float sc = a.f0 * 1.4426950408889634f;
        *reinterpret_cast<float4*>(qs + warp * D + lane * 4) = make_float4(f[0] * sc, f[1] * sc, f[2] * sc, f[3] * sc);
      } else {
        uint2 packed;
        packed.x = ptx::pack_bf16x2(f[0], f[1]);
        packed.y = ptx::pack_bf16x2(f[2], f[3]);
        bf16* dst = static_cast<bf16*>(const_cast<void*>(role == 1 ? a.in1 : a.in2)) + kv_off(pos) + lane * 4;
        *reinterpret_cast<uint2*>(dst) = packed;
      }
    }
  }
  float m[GM], l[GM], o[GM][4];
#pragma unroll
  for (int g = 0; g < GM; ++g) {
#pragma unroll
    for (int i = 0; i < 4; ++i) o[g][i] = 0.f;
    m[g] = -CUDART_INF_F;
    l[g] = 0.f;
  }
  __syncthreads();
  for (int p0 = p_first; p0 < s_end; p0 += NW * UN) {  // warp-uniform trip count
#pragma unroll
    for (int u = 0; u < UN; ++u) {
      if (p0 + u < s_end && !(p0 == p_first && pre[u])) {
        const size_t off = kv_off(p0 + u) + lane * 4;
        kr[u] = *reinterpret_cast<const uint2*>(kc + off);
        vr[u] = *reinterpret_cast<const uint2*>(vc + off);
      }
    }
#pragma unroll
    for (int u = 0; u < UN; ++u) {
      if (p0 + u >= s_end) break;
      const float2 k01 = ptx::unpack_bf16x2(kr[u].x), k23 = ptx::unpack_bf16x2(kr[u].y);
      const float2 v01 = ptx::unpack_bf16x2(vr[u].x), v23 = ptx::unpack_bf16x2(vr[u].y);
#pragma unroll
      for (int g = 0; g < GM; ++g) {
        if (g >= G) break;
        const float4 qa = *reinterpret_cast<const float4*>(qs + g * D + lane * 4);
        float sc = qa.x * k01.x;
        sc = fmaf(qa.y, k01.y, sc);
        sc = fmaf(qa.z, k23.x, sc);
        sc = fmaf(qa.w, k23.y, sc);
        sc = ptx::warp_reduce_sum(sc);
        const float mn = fmaxf(m[g], sc);
        const float corr = exp2f(m[g] - mn), pe = exp2f(sc - mn);
        l[g] = l[g] * corr + pe;
        o[g][0] = fmaf(pe, v01.x, o[g][0] * corr);
        o[g][1] = fmaf(pe, v01.y, o[g][1] * corr);
        o[g][2] = fmaf(pe, v23.x, o[g][2] * corr);
        o[g][3] = fmaf(pe, v23.y, o[g][3] * corr);
        m[g] = mn;
      }
    }
  }
  // reduce the 16 warps through shared memory: [NW][G] maxima and sums, [NW][G][D] outputs (66 KB for G = 8)
#pragma unroll
  for (int g = 0; g < GM; ++g) {
    if (g >= G) break;
    if (lane == 0) {
      sh_m[warp * G + g] = m[g];
      sh_l[warp * G + g] = l[g];
    }
    *reinterpret_cast<float4*>(sh_o + (warp * G + g) * D + lane * 4) = make_float4(o[g][0], o[g][1], o[g][2], o[g][3]);
  }
  __syncthreads();
  float* partial = static_cast<float*>(a.out0);
  for (int idx = threadIdx.x; idx < G * D; idx += kThreads) {
    const int g = idx / D, d = idx % D;
    float mm = -CUDART_INF_F;
    for (int r = 0; r < NW; ++r) mm = fmaxf(mm, sh_m[r * G + g]);
    float ll = 0.f, acc = 0.f;
    for (int r = 0; r < NW; ++r) {
      const float mr = sh_m[r * G + g];
      const float w = mr == -CUDART_INF_F ? 0.f : exp2f(mr - mm);
      ll += sh_l[r * G + g] * w;
      acc += sh_o[(r * G + g) * D + d] * w;
    }
    float* dst = partial + ((size_t(b) * Hq + h0 + g) * S + split) * (D + 1);
    dst[d] = ll > 0.f ? acc / ll : 0.f;
    if (d == 0) dst[D] = ll > 0.f ? (mm + log2f(ll)) * 0.6931471805599453f : -CUDART_INF_F;
  }
  __syncthreads();
  if (FUSED) {
    // arrival counter of this (batch, kv head): the split that brings it to step * S combines all S partials
    uint32_t* cnt = reinterpret_cast<uint32_t*>(partial + size_t(a.M) * Hq * S * (D + 1)) + (b * Hkv + kvh);
    __shared__ int s_last;
    if (threadIdx.x == 0) {
      __threadfence();
      s_last = atomicAdd(cnt, 1u) + 1u == step * uint32_t(S);
    }
    __syncthreads();
    if (s_last) {
      __threadfence();
      bf16* outp = static_cast<bf16*>(a.out1);
      // all loads of a (head, dim) pair are issued before any is used: one L2 round trip instead of 2 * S
      constexpr int kMaxS = 16;
      for (int idx = threadIdx.x; idx < G * D; idx += kThreads) {
        const int g = idx / D, d = idx % D;
        const float* base = partial + (size_t(b) * Hq + h0 + g) * S * (D + 1);
        float den = 0.f, acc = 0.f, mx = -CUDART_INF_F;
        for (int s0 = 0; s0 < S; s0 += kMaxS) {  // S <= 16 in practice: one pass
          float lse[kMaxS], val[kMaxS];
#pragma unroll
          for (int j = 0; j < kMaxS; ++j) {
            const bool on = s0 + j < S;
            lse[j] = on ? __ldcg(base + size_t(s0 + j) * (D + 1) + D) : -CUDART_INF_F;
            val[j] = on ? __ldcg(base + size_t(s0 + j) * (D + 1) + d) : 0.f;
          }
          float m2 = mx;
#pragma unroll
          for (int j = 0; j < kMaxS; ++j) m2 = fmaxf(m2, lse[j]);
          const float resc = mx == -CUDART_INF_F ? 0.f : __expf(mx - m2);
          den *= resc;
          acc *= resc;
#pragma unroll
          for (int j = 0; j < kMaxS; ++j) {
            const float w = lse[j] == -CUDART_INF_F ? 0.f : __expf(lse[j] - m2);
            den += w;
            acc += w * val[j];
          }
          mx = m2;
        }
        outp[(size_t(b) * Hq + h0 + g) * D + d] = __float2bfloat16(den > 0.f ? acc / den : 0.f);
      }
    }
    __syncthreads();
  }
}

// combine: in0 partial [B,Hq,S,129] -> out0 [B,Hq,128] bf16. tile = 4 (b,h) pairs.
__device__ __forceinline__ void attn_combine_body(const OpArgs& a, int tile) {
  constexpr int D = 128;
  const int S = a.i4, Hq = a.i0;
  const int bh = tile * (kThreads / D) + threadIdx.x / D, d = threadIdx.x % D;
  if (bh >= a.M * Hq) return;
  const float* base = static_cast<const float*>(a.in0) + size_t(bh) * S * (D + 1);
  float mx = -CUDART_INF_F;
  for (int s = 0; s < S; ++s) mx = fmaxf(mx, base[size_t(s) * (D + 1) + D]);
  float den = 0.f, acc = 0.f;
  for (int s = 0; s < S; ++s) {
    const float lse = base[size_t(s) * (D + 1) + D];
    const float w = lse == -CUDART_INF_F ? 0.f : __expf(lse - mx);
    den += w;
    acc += w * base[size_t(s) * (D + 1) + d];
  }
  static_cast<bf16*>(a.out0)[size_t(bh) * D + d] = __float2bfloat16(den > 0.f ? acc / den : 0.f);
}

// ------------------------------------------------------------------ all-reduce (+ residual add)
// in0: symmetric partial buffer (bf16, N elements per parity, i2 = parity stride bytes); in3: residual (updated in
// place: residual += sum) or null; out0: reduced output (bf16) or null when only the residual is wanted.
// i0 = AR op ordinal (flag row block). Per-tile cross-rank flag exchange, then NVLS ld_reduce (P2P sum without NVLS).
__device__ __forceinline__ void allreduce_body(const KernelParams& p, const OpArgs& a, int tile, uint32_t step) {
  const SymmCtx& c = p.ctx;
  const int W = c.world;
  uint32_t* flags = p.ar_flags + (size_t(a.i0) * kMaxArTiles + tile) * TDIST_MAX_RANKS;
  if (W > 1) barrier_all_block(c, flags, step);
  char* buf = static_cast<char*>(const_cast<void*>(a.in0)) + (a.i2 ? size_t(step & 1) * a.i2 : 0);
  const long long nv = (long long)a.N >> 3;
  const long long per = (nv + a.tiles - 1) / a.tiles;
  const long long beg = (long long)tile * per, end = min(beg + per, nv);
  bf16* res = static_cast<bf16*>(const_cast<void*>(a.in3));
  bf16* out = static_cast<bf16*>(a.out0);
  for (long long i = beg + threadIdx.x; i < end; i += kThreads) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (W > 1 && c.mc_base != nullptr) {
      unpack8(ptx::multimem_ld_reduce_bf16x8(mc_ptr(c, buf) + i * 16), acc);
    } else {
      for (int s = 0; s < W; ++s) {
        float part[8];
        unpack8(ptx::ld_v4(symm_at(c, buf, (c.rank + s) % W) + i * 16), part);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += part[j];
      }
    }
    if (res != nullptr) {
      // the summed projection is rounded to bf16 first, exactly like an unfused all-reduce followed by an add
      float r[8];
      unpack8(ptx::ld_v4(res + i * 8), r);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = __bfloat162float(__float2bfloat16(acc[j])) + r[j];
      *reinterpret_cast<int4*>(res + i * 8) = pack8(acc);
    }
    if (out != nullptr) *reinterpret_cast<int4*>(out + i * 8) = pack8(acc);
  }
}

__device__ __forceinline__ void prefetch_body(const OpArgs& a, int tile) {
  const size_t bytes = size_t(a.N) * a.K * 2;
  const size_t piece = 4096;
  const size_t pieces = (bytes + piece - 1) / piece;
  const char* base = static_cast<const char*>(a.in1);
  for (size_t i = size_t(tile) * kThreads + threadIdx.x; i < pieces; i += size_t(a.tiles) * kThreads) {
    const size_t off = i * piece;
    ptx::prefetch_l2_bulk(base + off, uint32_t(min(piece, bytes - off)));
  }
}

__device__ __forceinline__ void embedding_body(const OpArgs& a, int tile) {
  const long long id = static_cast<const long long*>(a.in0)[tile];
  const int4* src = reinterpret_cast<const int4*>(static_cast<const bf16*>(a.in1) + size_t(id) * a.K);
  int4* dst = reinterpret_cast<int4*>(static_cast<bf16*>(a.out0) + size_t(tile) * a.K);
  for (int i = threadIdx.x; i < a.K / 8; i += kThreads) dst[i] = src[i];
}

__device__ __forceinline__ void argmax_body(const OpArgs& a, int tile, uint8_t* smem) {
  float* sv = reinterpret_cast<float*>(smem);
  int* si = reinterpret_cast<int*>(smem + 64 * sizeof(float));
  const bf16* row = static_cast<const bf16*>(a.in0) + size_t(tile) * a.i1;
  float best = -CUDART_INF_F;
  int bi = 0;
  for (int i = threadIdx.x; i < a.N; i += kThreads) {
    const float v = __bfloat162float(row[i]);
    if (v > best) {
      best = v;
      bi = i;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) {
      best = ov;
      bi = oi;
    }
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) {
    sv[threadIdx.x >> 5] = best;
    si[threadIdx.x >> 5] = bi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (kThreads >> 5); ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) {
        best = sv[w];
        bi = si[w];
      }
    static_cast<long long*>(a.out0)[tile] = bi;
  }
}

// ------------------------------------------------------------------ the interpreter
// Task records and op arguments live in global memory; a decode step streams gigabytes of weights through L2, so by
// the time a CTA reaches task i + 1 both are usually cold (two dependent DRAM round trips on the critical path of
// every task). They are therefore prefetched one task ahead by 12 threads into a double-buffered shared-memory slot,
// published by the __syncthreads that ends every task.
struct TaskSlot {
  OpArgs a;   // 160 bytes = 10 x int4
  Task t;     // 32 bytes = 2 x int4
};
static_assert(sizeof(OpArgs) == 160 && sizeof(Task) == 32 && sizeof(TaskSlot) == 192, "TaskSlot layout");

__global__ void __launch_bounds__(kThreads, 1) megakernel(const __grid_constant__ KernelParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const PhaseState ps{p.step_state, p.step_state + 1};
  const uint32_t base = phase_begin(ps);
  const uint32_t step = base + 1;
  const int n = p.queue_len[blockIdx.x];
  const Task* q = p.queues + size_t(blockIdx.x) * p.max_tasks;
  // weight buffers (behind the per-task scratch area), their barriers, the task slots and the stream table
  Ring r;
  r.buf = smem + p.ring_offset;
  r.slots = uint32_t(p.ring_stages);
  r.full = reinterpret_cast<uint64_t*>(r.buf + size_t(r.slots) * kSlotBytes);
  TaskSlot* slots = reinterpret_cast<TaskSlot*>(r.buf + size_t(r.slots) * kSlotBytes + 256);
  StreamEntry* tab = reinterpret_cast<StreamEntry*>(slots + 2);
  const uint32_t n_ent = uint32_t(p.stream_len[blockIdx.x]);
  r.tab = tab;
  r.n_ent = n_ent;
  {
    const int4* src = reinterpret_cast<const int4*>(p.streams + size_t(blockIdx.x) * p.max_streams);
    int4* dst = reinterpret_cast<int4*>(tab);
    for (uint32_t i = threadIdx.x; i < n_ent * (sizeof(StreamEntry) / 16); i += kThreads) dst[i] = src[i];
  }
  int4 pf = make_int4(0, 0, 0, 0);
  int op_next = 0;  // op index of task i + 1 (threads 0..9)
  if (n > 0) {
    if (threadIdx.x < 10) {
      pf = reinterpret_cast<const int4*>(p.ops + q[0].op)[threadIdx.x];
      if (n > 1) op_next = q[1].op;
    } else if (threadIdx.x < 12) {
      pf = reinterpret_cast<const int4*>(q)[threadIdx.x - 10];
    }
    if (threadIdx.x < 12) reinterpret_cast<int4*>(&slots[0])[threadIdx.x] = pf;
  }
  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < r.slots; ++s) ptx::mbar_init(&r.full[s], 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  r.e_idx = 0;
  r.e_first = 0;
  r.e_total = n_ent > 0 ? tab[0].total : 0;
  r.uses = 0;
  r.issued = 0xffffffffu;
  r.prof = p.trace != nullptr;
  r.cyc[0] = r.cyc[1] = r.cyc[2] = r.cyc[3] = 0;
  r.pf_idx = 0;
  r.pf_first = 0;
  r.pf_total = r.e_total;
  r.pf_next = (threadIdx.x >> 5) + r.slots;  // (stage `warp` itself is loaded right away)
  if ((threadIdx.x >> 5) < r.slots) issue_stage(r, threadIdx.x >> 5);  // every buffer starts with its first stage
  uint32_t stage_base = 0;  // id of the first stage of the next linear task of this CTA
  uint32_t lin_k = 0;       // ordinal of the next linear task of this CTA (index into the stream table)
  // TMA descriptors live in global memory and are cold when an op is reached (a step streams GBs through L2); a
  // descriptor miss stalls the issuing warp, so they are prefetched a few linear tasks ahead
  if (threadIdx.x < 4 && threadIdx.x < n_ent) ptx::prefetch_tensormap(tab[threadIdx.x].tmap);
  uint64_t* trace = p.trace != nullptr ? p.trace + (size_t(blockIdx.x) * p.max_tasks) * kTraceWords : nullptr;
  for (int i = 0; i < n; ++i) {
    const Task t = slots[i & 1].t;
    const OpArgs& a = slots[i & 1].a;
    const uint32_t tgt0 = step * uint32_t(t.dep0_cnt), tgt1 = step * uint32_t(t.dep1_cnt);
    uint64_t* tr = trace != nullptr ? trace + size_t(i) * kTraceWords : nullptr;
    if (threadIdx.x == 0 && tr != nullptr) tr[0] = ptx::globaltimer();
    // start fetching task i + 1 (its op index was read one task earlier, so the two loads are not chained)
    int op_next2 = 0;
    if (i + 1 < n) {
      if (threadIdx.x < 10) {
        pf = reinterpret_cast<const int4*>(p.ops + op_next)[threadIdx.x];
        if (i + 2 < n) op_next2 = q[i + 2].op;
      } else if (threadIdx.x < 12) {
        pf = reinterpret_cast<const int4*>(q + i + 1)[threadIdx.x - 10];
      }
    }
    if (t.type == OP_LINEAR) {
      // waits for its dependencies inside (its first weight stages are already in flight)
      if (threadIdx.x >= 32 && threadIdx.x < 34 && lin_k + 3 + (threadIdx.x - 32) < n_ent)
        ptx::prefetch_tensormap(tab[lin_k + 3 + (threadIdx.x - 32)].tmap);
      linear_body(a, t.tile, smem, step, p.scoreboard, t, tgt0, tgt1, r, stage_base, lin_k, tr);
      ++lin_k;
    } else {
      if (t.type != OP_ATTN_SPLIT && t.type != OP_ATTN_FUSED) {  // (attention waits inside, after it has started its K/V loads)
        if (threadIdx.x == 0) {
          wait_slot(p.scoreboard, t.dep0, tgt0);
          wait_slot(p.scoreboard, t.dep1, tgt1);
          if (tr != nullptr) tr[2] = ptx::globaltimer();  // dependencies satisfied
        }
        __syncthreads();
      }
      switch (t.type) {
        case OP_RMS_NORM: rmsnorm_body(a, t.tile, smem); break;
        case OP_SILU_MUL: silu_mul_body(a, t.tile); break;
        case OP_ADD: add_body(a, t.tile); break;
        case OP_QK_NORM_ROPE_KV: qk_rope_body(a, t.tile); break;
        case OP_ATTN_SPLIT: {
          const int g = a.i0 / a.i1;
          for (int done = 0; done < g; done += 8)  // at most 8 query heads per pass over the split's K/V
            attn_split_body<false>(a, t.tile, smem, min(8, g - done), g, done, done == 0 ? p.scoreboard : nullptr, t, tgt0, tgt1, tr, step);
          break;
        }
        case OP_ATTN_FUSED:  // (the builder only emits this op for <= 8 query heads per KV head)
          attn_split_body<true>(a, t.tile, smem, a.i0 / a.i1, a.i0 / a.i1, 0, p.scoreboard, t, tgt0, tgt1, tr, step);
          break;
        case OP_ATTN_COMBINE: attn_combine_body(a, t.tile); break;
        case OP_ALLREDUCE: allreduce_body(p, a, t.tile, step); break;
        case OP_BARRIER:
          if (p.ctx.world > 1) barrier_all_block(p.ctx, p.ar_flags + (size_t(a.i0) * kMaxArTiles) * TDIST_MAX_RANKS, step);
          break;
        case OP_PREFETCH: prefetch_body(a, t.tile); break;
        case OP_EMBEDDING: embedding_body(a, t.tile); break;
        case OP_ARGMAX: argmax_body(a, t.tile, smem); break;
        default: break;
      }
    }
    if (i + 1 < n && threadIdx.x < 12) reinterpret_cast<int4*>(&slots[(i + 1) & 1])[threadIdx.x] = pf;
    op_next = op_next2;
    __syncthreads();
    if (threadIdx.x == 0) {
      if (t.out_slot >= 0) {
#ifdef TDIST_MEGA_SC_FENCE
        __threadfence();
        atomicAdd(p.scoreboard + t.out_slot, 1u);
#else
        // acq_rel fence + fire-and-forget reduction instead of __threadfence() (a sequentially consistent fence) +
        // atomicAdd (which waits for the old value): this thread signals hundreds of tasks per token
        ptx::fence_acq_rel_gpu();
        ptx::red_add_relaxed_gpu(p.scoreboard + t.out_slot, 1u);
#endif
      }
      if (tr != nullptr) tr[1] = ptx::globaltimer();
    }
  }
  if (r.prof && (threadIdx.x & 31) == 0) {
    // per-warp cycle totals behind the task stamps: [num_ctas][16 warps][4]
    uint64_t* w = p.trace + size_t(gridDim.x) * p.max_tasks * kTraceWords + (size_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5)) * 4;
    for (int j = 0; j < 4; ++j) w[j] = r.cyc[j];
  }
  phase_end(ps, base, 1);
}

}  // namespace mega
}  // namespace tdist

extern "C" int tdist_mega_launch(const void* params, int num_ctas, int smem_bytes, void* stream) {
  using namespace tdist::mega;
  static int attr_set = 0;
  if (attr_set < smem_bytes) {
    cudaError_t e = cudaFuncSetAttribute(megakernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) {
      tdist::set_error(std::string("megakernel smem attribute: ") + cudaGetErrorString(e));
      return -1;
    }
    attr_set = smem_bytes;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_ctas);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const KernelParams* kp = static_cast<const KernelParams*>(params);
  cudaError_t e = cudaLaunchKernelEx(&cfg, megakernel, *kp);
  if (e != cudaSuccess) {
    tdist::set_error(std::string("megakernel launch: ") + cudaGetErrorString(e));
    return -1;
  }
  return 0;
}
