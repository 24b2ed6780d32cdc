// mega.h — shared definitions of the persistent decode megakernel (SURVEY.md L9 / M*).
//
// Reference: python/triton_dist/mega_triton_kernel/{core,tasks,kernels,models}. There a Python ModelBuilder turns
// every op into per-tile Task records, a scheduler deals them round-robin into per-SM work queues, and a code
// generator emits ONE Triton kernel with an if/elif per task type (core/code_generator.py:31-89,
// core/scheduler.py:40-95, models/model_builder.py:83-404). Here the builder and scheduler are native C++
// (csrc/mega/builder.cc) and the kernel is a hand-written CUDA interpreter (csrc/mega/megakernel.cu).
#pragma once
#include <stdint.h>

#include "../device/symm.cuh"

namespace tdist {
namespace mega {

enum OpType : int {
  OP_LINEAR = 0,       // y = [rmsnorm(x)] @ W^T [; silu(gate)*up pairing] [+ bias]      (qkv_proj / o_proj / fc1 / fc2 / linear)
  OP_RMS_NORM = 1,     // y = rmsnorm(x [+ residual]) * w
  OP_SILU_MUL = 2,     // y = silu(x[:, :I]) * x[:, I:]
  OP_ADD = 3,          // y = a + b
  OP_QK_NORM_ROPE_KV = 4,
  OP_ATTN_SPLIT = 5,
  OP_ATTN_COMBINE = 6,
  OP_ALLREDUCE = 7,    // NVLS one-shot all-reduce of a symmetric buffer [+ residual add]
  OP_BARRIER = 8,      // cross-rank barrier
  OP_PREFETCH = 9,     // L2 prefetch of a weight matrix (no dependency)
  OP_EMBEDDING = 10,
  OP_ARGMAX = 11,
  OP_ATTN_FUSED = 12,  // qk-norm + RoPE + KV-cache write + split-KV attention + LSE combine in one task type
  OP_NUM_TYPES
};

// One record per op (NOT per tile); tasks reference it by index.
struct OpArgs {
  const void* in0;   // activations / query / lhs
  const void* in1;   // weight / key cache / rhs
  const void* in2;   // norm weight / value cache / bias
  const void* in3;   // residual / block table / q-norm weight
  const void* in4;   // kv_lens or positions / k-norm weight
  const void* in5;   // cos-sin cache
  void* out0;
  void* out1;
  void* out2;
  int M, N, K;       // rows (batch), output features, reduction dim
  int i0, i1, i2, i3, i4, i5;  // op-specific integers (heads, page size, strides, flags ...)
  long long l0, l1, l2;        // op-specific strides
  float f0, f1;                // eps, scale
  int tiles;                   // number of tiles of this op
  int cols_per_tile;           // LINEAR: output columns per tile
  int pad_[2];                 // sizeof == 160 (copied to shared memory as 10 x int4)
};

// Fixed-size task record (the reference's INT_PER_TASK words, core/scheduler.py:40-62).
struct Task {
  int op;        // index into the op table
  int type;      // OpType (duplicated for a cheap switch)
  int tile;      // tile id inside the op
  int dep0;      // scoreboard slot to wait on (-1 = none)
  int dep0_cnt;  // tiles that must have completed in that slot per step
  int dep1;
  int dep1_cnt;
  int out_slot;  // scoreboard slot this task increments when done
};

// Geometry of the weight stream of one LINEAR tile (identical on host and device): the tile's cols_per_tile rows of
// W are cut into nblk row blocks of R <= kStageRows rows, K into kboxes extents of kstage = 64 * nsub elements; one
// stage = one (row block, k extent) = one TMA box {64, R, nsub} of the 3-D view [k % 64][row][k / 64] of W.
constexpr int kStageRows = 24;
struct LinearGeom {
  int nblk, R, nsub, kstage, kboxes;
};
#ifdef __CUDACC__
__host__ __device__
#endif
inline LinearGeom linear_geom(int cols_per_tile, int K) {
  LinearGeom g;
  g.nblk = (cols_per_tile + kStageRows - 1) / kStageRows;
  g.R = (cols_per_tile + g.nblk - 1) / g.nblk;
  // a stage buffer holds kStageRows * 4 lines of 128 B, a stage nsub slabs of R lines (64 k-elements each): short row
  // blocks get proportionally longer k extents, so that a stage is always ~12 KB (4 slabs = 256 k for R = 24,
  // 12 slabs = 768 k for R = 8).
  g.nsub = (kStageRows * 4) / g.R;
  if (g.nsub > 12) g.nsub = 12;
  const int need = (K + 63) / 64;  // never more slabs than K has
  if (g.nsub > need) g.nsub = need;
  g.kstage = g.nsub * 64;
  g.kboxes = (K + g.kstage - 1) / g.kstage;
  return g;
}

// One entry per LINEAR task of a CTA's queue, in queue order. Stages are numbered consecutively over all linear
// tasks of the CTA, per task in the order (blob, row block, k box).
struct StreamEntry {
  const void* tmap;     // device copy of the op's CUtensorMap: W as [64][rows][K / 64] bf16, box {64, R, nsub}, 128 B swizzle
  uint32_t row0;        // first row of the tile (c_beg)
  uint32_t pair_rows;   // silu pairing: row offset of the "up" half (N), else 0
  uint8_t R, nblk;      // R <= kStageRows, nblk <= 11
  uint16_t kstage;      // k-elements per stage (64 * number of boxes)
  uint32_t kboxes;
  uint32_t first;       // id of the task's first stage
  uint32_t total;       // stages of the task = blobs * nblk * kboxes
};
static_assert(sizeof(StreamEntry) == 32, "StreamEntry is copied as 2 x int4");

struct KernelParams {
  const Task* queues;      // [num_ctas][max_tasks]
  const int* queue_len;    // [num_ctas]
  const OpArgs* ops;
  uint32_t* scoreboard;    // [num_slots], monotone: value after step s == s * tiles(op)
  uint32_t* step_state;    // [2]: step counter, done counter (PhaseState)
  uint32_t* ar_flags;      // symmetric: cross-rank flags [num_ar_ops * kMaxArTiles][TDIST_MAX_RANKS] + barrier rows
  SymmCtx ctx;
  int max_tasks;
  int ring_offset;         // byte offset of the weight ring in dynamic shared memory (behind every op's scratch area)
  int ring_stages;         // chunks in the ring
  const StreamEntry* streams;  // [num_ctas][max_streams]
  const int* stream_len;       // [num_ctas]
  int max_streams;
  uint64_t* trace;         // optional [num_ctas][max_tasks][kTraceWords] globaltimer stamps per task (nullptr = off)
};

constexpr int kThreads = 512;
constexpr int kMaxArTiles = 32;
constexpr int kTraceWords = 6;         // per task: begin, end, [linear: deps satisfied, activations staged, stream consumed], spare
constexpr int kSlotBytes = kStageRows * 256 * 2;   // stage buffer: up to 24 rows x 256 k-elements (four {64, R} TMA boxes)
constexpr int kMaxRingStages = 16;     // one stage buffer per warp at most
constexpr int kPrefetchRounds = 1;     // L2 prefetch distance behind every stage load, in buffer-rounds (slots stages); 0 = off
constexpr int kIdlePrefetchRounds = 0; // L2 prefetch distance while a CTA waits on the scoreboard; 0 = off (measured: no gain,
                                       // the stage loads are latency- not bandwidth-limited at TP8 tile sizes)
constexpr int kSmemLimit = 232448;     // 227 KB opt-in limit per CTA on sm_100

}  // namespace mega
}  // namespace tdist
