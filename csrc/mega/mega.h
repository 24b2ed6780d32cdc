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

struct KernelParams {
  const Task* queues;      // [num_ctas][max_tasks]
  const int* queue_len;    // [num_ctas]
  const OpArgs* ops;
  uint32_t* scoreboard;    // [num_slots], monotone: value after step s == s * tiles(op)
  uint32_t* step_state;    // [2]: step counter, done counter (PhaseState)
  uint32_t* ar_flags;      // symmetric: cross-rank flags [num_ar_ops * kMaxArTiles][TDIST_MAX_RANKS] + barrier rows
  SymmCtx ctx;
  int max_tasks;
};

constexpr int kThreads = 512;
constexpr int kMaxArTiles = 32;

}  // namespace mega
}  // namespace tdist
