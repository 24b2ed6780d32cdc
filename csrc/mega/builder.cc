// builder.cc — native task-graph builder + static scheduler of the decode megakernel.
//
// Counterpart of the reference's Python ModelBuilder/TaskBuilder/Scheduler
// (mega_triton_kernel/models/model_builder.py:83-404, core/builder.py, core/scheduler.py:40-95): every op is
// expanded into per-tile tasks, dependencies default to "all tiles of the previous op" (model_builder.py:141-146),
// tasks are dealt round-robin into one queue per CTA, and the queues + op table are uploaded once. run() is a single
// cooperative launch of the interpreter kernel.
#include <cuda_runtime.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/tdist.h"
#include "../runtime/driver.h"
#include "mega.h"

extern "C" int tdist_mega_launch(const void* params, int num_ctas, int smem_bytes, void* stream);

namespace tdist {
namespace mega {

struct Builder {
  int rank = 0, world = 1, num_ctas = 148;
  std::vector<OpArgs> ops;
  std::vector<int> types, dep0, dep1;
  std::vector<bool> chain;  // op takes part in the default dependency chain
  int num_ar_ops = 0;
  // compiled state
  bool compiled = false;
  Task* d_queues = nullptr;
  int* d_queue_len = nullptr;
  OpArgs* d_ops = nullptr;
  uint32_t* d_scoreboard = nullptr;
  uint32_t* d_step = nullptr;
  KernelParams params;
  int max_tasks = 0, total_tasks = 0, smem_bytes = 0;
};

static int tiles_of(const OpArgs& a, int type, int num_ctas, int* cols_per_tile) {
  *cols_per_tile = 0;
  switch (type) {
    case OP_LINEAR: {
      // columns per tile: enough tiles to cover every CTA ~2x when the matrix is big, never less than 16 columns
      // (one per warp) so that each warp streams whole rows
      int cpt = (a.N + 2 * num_ctas - 1) / (2 * num_ctas);
      cpt = std::max(16, (cpt + 15) / 16 * 16);
      *cols_per_tile = cpt;
      return (a.N + cpt - 1) / cpt;
    }
    case OP_RMS_NORM: return a.M;
    case OP_SILU_MUL: return std::max(1, std::min(num_ctas, int(((long long)a.M * a.N / 8 + kThreads - 1) / kThreads)));
    case OP_ADD: return std::max(1, std::min(num_ctas, int((a.N / 8 + kThreads - 1) / kThreads)));
    case OP_QK_NORM_ROPE_KV: return std::max(1, std::min(num_ctas, (a.M * (a.i0 + 2 * a.i1) + 15) / 16));
    case OP_ATTN_SPLIT: return a.M * a.i1 * a.i4;
    case OP_ATTN_COMBINE: return (a.M * a.i0 + 3) / 4;
    case OP_ALLREDUCE: return std::max(1, std::min(kMaxArTiles, int(((long long)a.N / 8 + 2 * kThreads - 1) / (2 * kThreads))));
    case OP_BARRIER: return 1;
    case OP_PREFETCH: return std::max(1, std::min(num_ctas, int(((size_t)a.N * a.K * 2 / 4096 + kThreads - 1) / kThreads)));
    case OP_EMBEDDING: return a.M;
    case OP_ARGMAX: return a.M;
  }
  return 1;
}

static int smem_of(const OpArgs& a, int type) {
  switch (type) {
    case OP_LINEAR: return a.M * a.K * 2 + 256;
    case OP_RMS_NORM: return a.K * 4 + 256;
    case OP_ATTN_SPLIT: {
      const int g = std::min(8, a.i0 / std::max(a.i1, 1));
      return (kThreads / 16) * g * (2 + 128) * 4;
    }
    case OP_ARGMAX: return 1024;
  }
  return 0;
}

}  // namespace mega
}  // namespace tdist

using namespace tdist;
using namespace tdist::mega;

extern "C" {

int tdist_mega_create(int rank, int world, int num_ctas, void** out) {
  Builder* b = new Builder();
  b->rank = rank;
  b->world = world;
  if (num_ctas <= 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_ctas, cudaDevAttrMultiProcessorCount, dev);
    if (num_ctas <= 0) num_ctas = 148;
  }
  b->num_ctas = num_ctas;
  *out = b;
  return 0;
}

// dep0/dep1: -2 = previous op in the chain (dep0 only), -1 = none, >= 0 = explicit op index. chain = 0 for ops that
// neither wait nor are waited upon by default (prefetch).
int tdist_mega_add_op(void* h, int type, const tdist_mega_op* d, int dep0, int dep1, int chain) {
  Builder* b = static_cast<Builder*>(h);
  if (b->compiled) {
    set_error("mega: builder already compiled");
    return -1;
  }
  if (type < 0 || type >= OP_NUM_TYPES) {
    set_error("mega: unknown op type");
    return -1;
  }
  OpArgs a;
  memset(&a, 0, sizeof(a));
  a.in0 = d->in[0]; a.in1 = d->in[1]; a.in2 = d->in[2]; a.in3 = d->in[3]; a.in4 = d->in[4]; a.in5 = d->in[5];
  a.out0 = d->out[0]; a.out1 = d->out[1]; a.out2 = d->out[2];
  a.M = d->M; a.N = d->N; a.K = d->K;
  a.i0 = d->i[0]; a.i1 = d->i[1]; a.i2 = d->i[2]; a.i3 = d->i[3]; a.i4 = d->i[4]; a.i5 = d->i[5];
  a.l0 = d->l[0]; a.l1 = d->l[1]; a.l2 = d->l[2];
  a.f0 = d->f[0]; a.f1 = d->f[1];
  if (type == OP_LINEAR && (a.M < 1 || a.M > 8 || a.K % 8 != 0)) {
    set_error("mega linear: need 1 <= M <= 8 rows and K % 8 == 0");
    return -1;
  }
  if (type == OP_ALLREDUCE || type == OP_BARRIER) a.i0 = b->num_ar_ops++;
  int cpt = 0;
  a.tiles = tiles_of(a, type, b->num_ctas, &cpt);
  a.cols_per_tile = cpt;
  const int idx = int(b->ops.size());
  if (dep0 == -2) {
    dep0 = -1;
    for (int i = idx - 1; i >= 0; --i)
      if (b->chain[i]) {
        dep0 = i;
        break;
      }
  }
  b->ops.push_back(a);
  b->types.push_back(type);
  b->dep0.push_back(dep0);
  b->dep1.push_back(dep1);
  b->chain.push_back(chain != 0);
  return idx;
}

int tdist_mega_compile(void* h, const tdist_symm_ctx* ctx, uint32_t* ar_flags) {
  Builder* b = static_cast<Builder*>(h);
  const int n_ops = int(b->ops.size());
  std::vector<std::vector<Task>> queues(b->num_ctas);
  int cursor = 0, smem = 1024;
  for (int o = 0; o < n_ops; ++o) {
    const OpArgs& a = b->ops[o];
    smem = std::max(smem, smem_of(a, b->types[o]));
    for (int t = 0; t < a.tiles; ++t) {
      Task k;
      k.op = o;
      k.type = b->types[o];
      k.tile = t;
      k.dep0 = b->dep0[o];
      k.dep0_cnt = k.dep0 >= 0 ? b->ops[k.dep0].tiles : 0;
      k.dep1 = b->dep1[o];
      k.dep1_cnt = k.dep1 >= 0 ? b->ops[k.dep1].tiles : 0;
      k.out_slot = o;
      queues[cursor % b->num_ctas].push_back(k);
      ++cursor;
    }
  }
  if (smem > 200 * 1024) {
    set_error("mega: an op needs more than 200 KB of shared memory (batch * K too large)");
    return -1;
  }
  b->smem_bytes = smem;
  b->total_tasks = cursor;
  b->max_tasks = 1;
  for (auto& q : queues) b->max_tasks = std::max(b->max_tasks, int(q.size()));
  std::vector<Task> flat(size_t(b->num_ctas) * b->max_tasks);
  std::vector<int> lens(b->num_ctas);
  memset(flat.data(), 0, flat.size() * sizeof(Task));
  for (int c = 0; c < b->num_ctas; ++c) {
    lens[c] = int(queues[c].size());
    std::copy(queues[c].begin(), queues[c].end(), flat.begin() + size_t(c) * b->max_tasks);
  }
  TDIST_RT_TRY(cudaMalloc(&b->d_queues, flat.size() * sizeof(Task)));
  TDIST_RT_TRY(cudaMalloc(&b->d_queue_len, lens.size() * sizeof(int)));
  TDIST_RT_TRY(cudaMalloc(&b->d_ops, b->ops.size() * sizeof(OpArgs)));
  TDIST_RT_TRY(cudaMalloc(&b->d_scoreboard, std::max(n_ops, 1) * sizeof(uint32_t)));
  TDIST_RT_TRY(cudaMalloc(&b->d_step, 2 * sizeof(uint32_t)));
  TDIST_RT_TRY(cudaMemcpy(b->d_queues, flat.data(), flat.size() * sizeof(Task), cudaMemcpyHostToDevice));
  TDIST_RT_TRY(cudaMemcpy(b->d_queue_len, lens.data(), lens.size() * sizeof(int), cudaMemcpyHostToDevice));
  TDIST_RT_TRY(cudaMemcpy(b->d_ops, b->ops.data(), b->ops.size() * sizeof(OpArgs), cudaMemcpyHostToDevice));
  TDIST_RT_TRY(cudaMemset(b->d_scoreboard, 0, std::max(n_ops, 1) * sizeof(uint32_t)));
  TDIST_RT_TRY(cudaMemset(b->d_step, 0, 2 * sizeof(uint32_t)));
  memset(&b->params, 0, sizeof(b->params));
  b->params.queues = b->d_queues;
  b->params.queue_len = b->d_queue_len;
  b->params.ops = b->d_ops;
  b->params.scoreboard = b->d_scoreboard;
  b->params.step_state = b->d_step;
  b->params.ar_flags = ar_flags;
  b->params.max_tasks = b->max_tasks;
  if (ctx != nullptr) {
    for (int i = 0; i < TDIST_MAX_RANKS; ++i) b->params.ctx.base[i] = ctx->base[i];
    b->params.ctx.mc_base = ctx->mc_base;
    b->params.ctx.bytes = ctx->bytes;
    b->params.ctx.rank = ctx->rank;
    b->params.ctx.world = ctx->world;
  } else {
    b->params.ctx.world = 1;
  }
  b->compiled = true;
  return 0;
}

int tdist_mega_run(void* h, void* stream) {
  Builder* b = static_cast<Builder*>(h);
  if (!b->compiled) {
    set_error("mega: compile() first");
    return -1;
  }
  return tdist_mega_launch(&b->params, b->num_ctas, b->smem_bytes, stream);
}

int tdist_mega_info(void* h, int* out6) {
  Builder* b = static_cast<Builder*>(h);
  out6[0] = int(b->ops.size());
  out6[1] = b->total_tasks;
  out6[2] = b->max_tasks;
  out6[3] = b->smem_bytes;
  out6[4] = b->num_ctas;
  out6[5] = b->num_ar_ops;
  return 0;
}

// tiles of op `idx` (after add_op) — lets the Python side size the symmetric flag area and check schedules
int tdist_mega_op_tiles(void* h, int idx) {
  Builder* b = static_cast<Builder*>(h);
  if (idx < 0 || idx >= int(b->ops.size())) return -1;
  return b->ops[idx].tiles;
}

// host copy of the schedule for inspection / CPU tests: writes up to cap tasks of queue `cta` as 8 ints each
int tdist_mega_get_queue(void* h, int cta, int* out, int cap) {
  Builder* b = static_cast<Builder*>(h);
  if (!b->compiled || cta < 0 || cta >= b->num_ctas) return -1;
  std::vector<Task> q(b->max_tasks);
  if (cudaMemcpy(q.data(), b->d_queues + size_t(cta) * b->max_tasks, q.size() * sizeof(Task), cudaMemcpyDeviceToHost) != cudaSuccess)
    return -1;
  int len = 0;
  cudaMemcpy(&len, b->d_queue_len + cta, sizeof(int), cudaMemcpyDeviceToHost);
  const int n = std::min(len, cap);
  for (int i = 0; i < n; ++i) memcpy(out + 8 * i, &q[i], sizeof(Task));
  return len;
}

int tdist_mega_destroy(void* h) {
  Builder* b = static_cast<Builder*>(h);
  if (b->compiled) {
    cudaDeviceSynchronize();
    cudaFree(b->d_queues);
    cudaFree(b->d_queue_len);
    cudaFree(b->d_ops);
    cudaFree(b->d_scoreboard);
    cudaFree(b->d_step);
  }
  delete b;
  return 0;
}

}  // extern "C"
