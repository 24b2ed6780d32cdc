// builder.cc — native task-graph builder + static scheduler of the decode megakernel.
//
// Counterpart of the reference's Python ModelBuilder/TaskBuilder/Scheduler
// (mega_triton_kernel/models/model_builder.py:83-404, core/builder.py, core/scheduler.py:40-95): every op is
// expanded into per-tile tasks, dependencies default to "all tiles of the previous op" (model_builder.py:141-146),
// tasks are dealt round-robin into one queue per CTA, and the queues + op table are uploaded once. run() is a single
// cooperative launch of the interpreter kernel.
#include <cuda_runtime.h>
#include <stdint.h>
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
  int max_tasks = 0, total_tasks = 0, smem_bytes = 0, ring_offset = 0, ring_stages = 0;
  uint64_t* d_trace = nullptr;
  StreamEntry* d_streams = nullptr;
  void* d_tmaps = nullptr;
  int* d_stream_len = nullptr;
  int max_streams = 1;
};

static int tiles_of(const OpArgs& a, int type, int num_ctas, int* cols_per_tile) {
  *cols_per_tile = 0;
  switch (type) {
    case OP_LINEAR: {
      // The kernel spreads a tile's weights evenly over all threads whatever its shape, so the only thing that
      // matters is that every CTA gets the same number of bytes: one tile per CTA per wave, at most 256 columns per
      // tile (bounds the per-tile column-sum array), at least 4 columns (one 16 KB chunk is ~1-4 rows).
      const int waves = std::max(1, (a.N + num_ctas * 256 - 1) / (num_ctas * 256));
      int tiles = std::min(waves * num_ctas, std::max(1, a.N / 4));
      int cpt = (a.N + tiles - 1) / tiles;
      *cols_per_tile = cpt;
      return (a.N + cpt - 1) / cpt;
    }
    case OP_RMS_NORM: return a.M;
    case OP_SILU_MUL: return std::max(1, std::min(num_ctas, int(((long long)a.M * a.N / 8 + kThreads - 1) / kThreads)));
    case OP_ADD: return std::max(1, std::min(num_ctas, int((a.N / 8 + kThreads - 1) / kThreads)));
    case OP_QK_NORM_ROPE_KV: return std::max(1, std::min(num_ctas, (a.M * (a.i0 + 2 * a.i1) + 15) / 16));
    case OP_ATTN_SPLIT:
    case OP_ATTN_FUSED: return a.M * a.i1 * a.i4;
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
    case OP_LINEAR: {
      // activations [M][kboxes * kstage + 8] bf16, norm weight [K] bf16 when fused, block sums [2][nblk][32][M] fp32,
      // cross-warp partials [2][16 warps][32][M] fp32, block_sum scratch
      const LinearGeom ge = linear_geom(a.cols_per_tile, a.K);
      const int xpitch = ge.kboxes * ge.kstage + 8;
      return a.M * xpitch * 2 + (a.in2 != nullptr ? a.K * 2 : 0) + 2 * ge.nblk * 32 * a.M * 4 + 2 * (kThreads / 32) * 32 * a.M * 4 + 256;
    }
    case OP_RMS_NORM: return a.K * 4 + 256;
    case OP_ATTN_SPLIT:
    case OP_ATTN_FUSED: {
      const int g = std::min(8, a.i0 / std::max(a.i1, 1));
      return (kThreads / 32) * g * (2 + 128) * 4 + g * 128 * 4;  // per-warp (max, sum, out[128]) for g heads + scaled q
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
  if (type == OP_LINEAR && (a.M < 1 || a.M > 8 || a.K % 64 != 0)) {
    set_error("mega linear: need 1 <= M <= 8 rows and K % 64 == 0");
    return -1;
  }
  if (type == OP_ATTN_FUSED) {
    const int g = a.i1 > 0 ? a.i0 / a.i1 : 0;
    if (g < 1 || g > 8 || a.i0 % std::max(a.i1, 1) != 0) {
      set_error("mega fused attention: needs 1..8 query heads per KV head");
      return -1;
    }
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
  // static weight-stream table: one entry per linear task of every CTA, in queue order (see mega.h StreamEntry), and
  // one TMA descriptor per linear op
  std::vector<int> tmap_of(n_ops, -1);
  std::vector<CUtensorMap> tmaps;
  {
    const DriverApi& d = driver();
    for (int o = 0; o < n_ops; ++o) {
      if (b->types[o] != OP_LINEAR) continue;
      const OpArgs& a = b->ops[o];
      if (!d.ok) {
        set_error("CUDA driver not available: " + d.err);
        return -1;
      }
      const LinearGeom ge = linear_geom(a.cols_per_tile, a.K);
      const uint64_t rows = (a.i0 & 1) ? uint64_t(a.N) * 2 : uint64_t(a.N);
      // 3-D view of the row-major [rows, K] matrix: dim0 = k % 64 (contiguous), dim1 = row (stride K * 2 bytes),
      // dim2 = k / 64 (stride 128 bytes). A box {64, R, nsub} lands in shared memory as [slab][row][128 B].
      cuuint64_t dims[3] = {64, rows, cuuint64_t(a.K / 64)};
      cuuint64_t strides[2] = {cuuint64_t(a.K) * 2, 128};
      cuuint32_t box[3] = {64u, cuuint32_t(ge.R), cuuint32_t(ge.nsub)};
      cuuint32_t estr[3] = {1, 1, 1};
      CUtensorMap tm;
      CUresult rc = d.cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(a.in1), dims, strides,
                                             box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (rc != CUDA_SUCCESS) {
        set_error("mega: cuTensorMapEncodeTiled failed for a linear weight (needs a 16-byte aligned [N, K] bf16 matrix): " +
                  cu_err(rc));
        return -1;
      }
      tmap_of[o] = int(tmaps.size());
      tmaps.push_back(tm);
    }
  }
  if (!tmaps.empty()) {
    TDIST_RT_TRY(cudaMalloc(&b->d_tmaps, tmaps.size() * sizeof(CUtensorMap)));
    TDIST_RT_TRY(cudaMemcpy(b->d_tmaps, tmaps.data(), tmaps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  }
  std::vector<std::vector<StreamEntry>> streams(b->num_ctas);
  int max_streams = 1;
  for (int c = 0; c < b->num_ctas; ++c) {
    uint32_t first = 0;
    for (const Task& k : queues[c]) {
      if (k.type != OP_LINEAR) continue;
      const OpArgs& a = b->ops[k.op];
      const LinearGeom ge = linear_geom(a.cols_per_tile, a.K);
      StreamEntry e;
      memset(&e, 0, sizeof(e));
      e.tmap = static_cast<const CUtensorMap*>(b->d_tmaps) + tmap_of[k.op];
      e.row0 = uint32_t(k.tile) * uint32_t(a.cols_per_tile);
      e.pair_rows = (a.i0 & 1) ? uint32_t(a.N) : 0u;
      e.R = uint8_t(ge.R);
      e.nblk = uint8_t(ge.nblk);
      e.kstage = uint16_t(ge.kstage);
      e.kboxes = uint32_t(ge.kboxes);
      e.total = ((a.i0 & 1) ? 2u : 1u) * uint32_t(ge.nblk) * uint32_t(ge.kboxes);
      e.first = first;
      first += e.total;
      streams[c].push_back(e);
    }
    max_streams = std::max(max_streams, int(streams[c].size()));
  }
  // dynamic shared memory = [per-task scratch, max over ops][weight ring][ring barriers][2 task slots][stream table]
  // (+4 KB: ldmatrix always reads 16 / 32 lines per slab, i.e. up to 31 lines past the last slab of a buffer; those
  // accumulator rows are never used, but the read must stay inside the allocation)
  const int aux_bytes = 256 + 2 * 192 + max_streams * int(sizeof(StreamEntry)) + 4096;
  const int ring_offset = (smem + 1023) / 1024 * 1024;  // 1024 B: the 128 B TMA swizzle is a function of the address
  int stages = (kSmemLimit - 2048 - ring_offset - aux_bytes) / kSlotBytes;  // 2 KB: static shared memory of the task bodies
  stages = std::min(stages, kMaxRingStages);
  if (stages < 2) {
    set_error("mega: an op needs too much shared memory to leave room for the weight ring (batch * K too large)");
    return -1;
  }
  b->ring_offset = ring_offset;
  b->ring_stages = stages;
  b->smem_bytes = ring_offset + stages * kSlotBytes + aux_bytes;
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
  std::vector<StreamEntry> sflat(size_t(b->num_ctas) * max_streams);
  std::vector<int> slens(b->num_ctas);
  memset(static_cast<void*>(sflat.data()), 0, sflat.size() * sizeof(StreamEntry));
  for (int c = 0; c < b->num_ctas; ++c) {
    slens[c] = int(streams[c].size());
    std::copy(streams[c].begin(), streams[c].end(), sflat.begin() + size_t(c) * max_streams);
  }
  TDIST_RT_TRY(cudaMalloc(&b->d_streams, sflat.size() * sizeof(StreamEntry)));
  TDIST_RT_TRY(cudaMalloc(&b->d_stream_len, slens.size() * sizeof(int)));
  TDIST_RT_TRY(cudaMemcpy(b->d_streams, sflat.data(), sflat.size() * sizeof(StreamEntry), cudaMemcpyHostToDevice));
  TDIST_RT_TRY(cudaMemcpy(b->d_stream_len, slens.data(), slens.size() * sizeof(int), cudaMemcpyHostToDevice));
  b->max_streams = max_streams;
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
  b->params.ring_offset = b->ring_offset;
  b->params.ring_stages = b->ring_stages;
  b->params.trace = nullptr;
  b->params.streams = b->d_streams;
  b->params.stream_len = b->d_stream_len;
  b->params.max_streams = b->max_streams;
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

// Per-task timeline (the counterpart of the reference's intra-kernel profiler): when enabled, thread 0 of every CTA
// stamps %globaltimer at the start and end of each task of the NEXT launches. get_trace copies the stamps of the last
// launch to the host as [num_ctas][max_tasks][6] nanosecond values: begin, end, and for linear tasks dependencies
// satisfied / activations staged / weight stream consumed (0 = not stamped).
int tdist_mega_set_trace(void* h, int enable) {
  Builder* b = static_cast<Builder*>(h);
  if (!b->compiled) {
    set_error("mega: compile() first");
    return -1;
  }
  if (enable && b->d_trace == nullptr) {
    const size_t bytes = (size_t(b->num_ctas) * b->max_tasks * kTraceWords + size_t(b->num_ctas) * (kThreads / 32) * 4) * sizeof(uint64_t);
    TDIST_RT_TRY(cudaMalloc(&b->d_trace, bytes));
    TDIST_RT_TRY(cudaMemset(b->d_trace, 0, bytes));
  }
  b->params.trace = enable ? b->d_trace : nullptr;
  return 0;
}

int tdist_mega_get_trace(void* h, unsigned long long* out, size_t cap_words) {
  Builder* b = static_cast<Builder*>(h);
  if (!b->compiled || b->d_trace == nullptr) {
    set_error("mega: tracing was not enabled");
    return -1;
  }
  const size_t words = size_t(b->num_ctas) * b->max_tasks * kTraceWords + size_t(b->num_ctas) * (kThreads / 32) * 4;
  if (cap_words < words) {
    set_error("mega: trace buffer too small");
    return -1;
  }
  TDIST_RT_TRY(cudaMemcpy(out, b->d_trace, words * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  return 0;
}

// Host-only planning query (no CUDA call, usable without a GPU): how a linear [N, K] (optionally gate|up paired) with
// batch M is cut for `num_ctas` CTAs. out12 = cols_per_tile, tiles, nblk, R, nsub, kstage, kboxes, stages per tile,
// scratch bytes, stage buffers that fit next to that scratch, stage buffer bytes, shared-memory limit.
int tdist_mega_linear_plan(int M, int N, int K, int pair, int has_norm, int num_ctas, int* out12) {
  if (M < 1 || M > 8 || N < 1 || K < 64 || K % 64 != 0 || num_ctas < 1) {
    set_error("mega_linear_plan: need 1 <= M <= 8, K a positive multiple of 64");
    return -1;
  }
  OpArgs a;
  memset(&a, 0, sizeof(a));
  a.M = M;
  a.N = N;
  a.K = K;
  a.i0 = pair ? 1 : 0;
  a.in2 = has_norm ? reinterpret_cast<const void*>(uintptr_t(16)) : nullptr;  // only tested for null-ness
  int cpt = 0;
  const int tiles = tiles_of(a, OP_LINEAR, num_ctas, &cpt);
  a.cols_per_tile = cpt;
  const LinearGeom ge = linear_geom(cpt, K);
  const int scratch = smem_of(a, OP_LINEAR);
  const int ring_offset = (scratch + 1023) / 1024 * 1024;
  int bufs = (kSmemLimit - 2048 - ring_offset - (256 + 2 * 192 + 4096)) / kSlotBytes;
  bufs = std::max(0, std::min(bufs, kMaxRingStages));
  const int v[12] = {cpt, tiles, ge.nblk, ge.R, ge.nsub, ge.kstage, ge.kboxes, (pair ? 2 : 1) * ge.nblk * ge.kboxes,
                     scratch, bufs, kSlotBytes, kSmemLimit};
  memcpy(out12, v, sizeof(v));
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
    if (b->d_trace) cudaFree(b->d_trace);
    cudaFree(b->d_streams);
    if (b->d_tmaps) cudaFree(b->d_tmaps);
    cudaFree(b->d_stream_len);
  }
  delete b;
  return 0;
}

}  // extern "C"
