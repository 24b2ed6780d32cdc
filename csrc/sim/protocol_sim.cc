// protocol_sim.cc — runs the tutorial-size distributed kernels of this library on the CPU, one OS thread per CTA (its CUDA
// threads are fibers) and one "GPU" per rank inside ONE process, under ThreadSanitizer. (SURVEY.md section 5 asks for a flag-protocol
// checker; the reference has none — its substitute is noise injection on real GPUs.)
//
// The kernel sources are compiled UNMODIFIED up to two mechanical steps done by tests/test_cpu_protocol_sim.py when it
// stages them: the files are cut before their launch wrappers (which use <<< >>>), and `__shared__ T name[..];` becomes a
// reference into a per-CTA buffer. csrc/sim/shim/kernels/common.cuh stands in for the device headers: flags become C++
// atomics with the memory orders the device code uses, payload accesses stay plain, __syncthreads() yields between fibers.
// ThreadSanitizer then reports any payload access that is not ordered by the flag protocol, a lost wake-up shows up as the
// 60 s wait bound, and the results are compared with a host computation. Ranks issue their calls independently (no
// barrier between calls), as streams on different GPUs do. TDIST_SIM_SLOW_RANK=r (+ TDIST_SIM_SLOW_MS) delays every flag write of rank r:
// the skew that opens the windows a protocol must survive.
//
//   protocol_sim queue   <ranks> <items> <item_vecs> <slots> <calls>
//   protocol_sim ag_gemm <ranks> <rows> <N> <K> <push_ctas> <ctas> <calls> [alternate: 1 = odd calls use rows / div, 2 = even calls] [div = 2]
//   protocol_sim gemm_rs <ranks> <M> <N> <Kl> <ctas> <calls> [alternate: 1 = odd calls use M / div, 2 = even calls] [div = 2]
#include "sim_runtime.h"

#include "tutorials/notify_wait_tutorial.cu"  // tdist::tut::tut_queue_kernel
#include "tutorials/overlap_tutorial.cu"      // tdist::tut::tut_ag_gemm_kernel, tut_gemm_rs_kernel

#include <atomic>
#include <string>

namespace {

using bf16 = __nv_bfloat16;
std::atomic<int> g_errors{0};

void mismatch(const char* what, int rank, int call, long i, float got, float want) {
  if (g_errors.fetch_add(1) < 8) fprintf(stderr, "%s: rank %d call %d element %ld: got %g, want %g\n", what, rank, call, i, got, want);
}

// ------------------------------------------------------------------ tutorial 1: producer / consumer queue
int run_queue(int world, int n_items, int item_vecs, int slots, int calls) {
  const size_t q_bytes = size_t(slots) * item_vecs * sizeof(int4);
  sim::Node node(world, q_bytes + 2 * slots * sizeof(uint32_t));
  auto value = [](int call, int rank, int item, int vec) { return make_int4(call, rank, item, vec); };
  std::vector<std::vector<int4>> inputs(size_t(calls) * world);  // one input per (call, rank): never rewritten
  for (int c = 0; c < calls; ++c)
    for (int r = 0; r < world; ++r) {
      auto& in = inputs[size_t(c) * world + r];
      in.resize(size_t(n_items) * item_vecs);
      for (int i = 0; i < n_items; ++i)
        for (int v = 0; v < item_vecs; ++v) in[size_t(i) * item_vecs + v] = value(c, r, i, v);
    }
  sim::run_ranks(world, [&](int rank) {
    const int prv = (rank + world - 1) % world;
    const tdist::SymmCtx ctx = node.ctx(rank);
    int4* queue = reinterpret_cast<int4*>(node.base(rank));
    uint32_t* full = reinterpret_cast<uint32_t*>(node.base(rank) + q_bytes);
    uint32_t* empty = full + slots;
    std::vector<int4> out(size_t(n_items) * item_vecs);
    uint32_t base = 0;
    for (int call = 0; call < calls; ++call) {
      std::memset(out.data(), 0xff, out.size() * sizeof(int4));
      const int4* in = inputs[size_t(call) * world + rank].data();
      int4* o = out.data();
      sim::launch(rank, 2, 256, [=] { tdist::tut::tut_queue_kernel(ctx, in, o, queue, full, empty, n_items, item_vecs, slots, base); });
      base += (n_items + slots - 1) / slots;
      for (int i = 0; i < n_items; ++i)
        for (int v = 0; v < item_vecs; ++v) {
          const int4 got = out[size_t(i) * item_vecs + v], want = value(call, prv, i, v);
          if (got.x != want.x || got.y != want.y || got.z != want.z || got.w != want.w)
            mismatch("queue", rank, call, long(i) * item_vecs + v, float(got.z), float(want.z));
        }
    }
  });
  printf("protocol_sim queue: %d ranks x %d calls x %d items (%d vecs) through %d slots: %s\n", world, calls, n_items, item_vecs,
         slots, g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

std::vector<bf16> random_bf16(size_t n, unsigned seed, float scale) {
  std::vector<bf16> v(n);
  for (auto& x : v) x = __float2bfloat16(sim::unit_rand(seed) * scale);
  return v;
}

// ------------------------------------------------------------------ tutorial 7: AllGather + GEMM
// Rank r: C_r[W * rows, N] = all_gather(A)[W * rows, K] * B_r[N, K]^T. Inputs differ per call; ranks do not wait for each
// other between calls, so a fast rank's call n+1 pushes into the OTHER half of a slow rank's workspace.
int run_ag_gemm(int world, int max_rows, int N, int K, int push_ctas, int ctas, int calls, int alternate, int alt_div) {
  const int rows_full = max_rows;
  const int blocks = (max_rows + 63) / 64;
  const size_t ws_bytes = size_t(2) * world * max_rows * K * sizeof(bf16);
  const size_t flag_bytes = size_t(world) * blocks * sizeof(uint32_t);
  sim::Node node(world, ws_bytes + flag_bytes + 64);
  std::vector<std::vector<bf16>> A(size_t(calls) * world), B(world);
  // --alternate: odd calls gather half as many rows on the same workspace and flags (prefill, then a shorter chunk)
  auto rows_of = [=](int call) { return alternate && (call & 1) == (alternate & 1) ? rows_full / alt_div : rows_full; };
  for (int r = 0; r < world; ++r) B[r] = random_bf16(size_t(N) * K, 77 + r, 0.25f);
  for (int c = 0; c < calls; ++c)
    for (int r = 0; r < world; ++r) A[size_t(c) * world + r] = random_bf16(size_t(rows_of(c)) * K, 1000 * c + r + 1, 1.f);
  sim::run_ranks(world, [&](int rank) {
    const tdist::SymmCtx ctx = node.ctx(rank);
    bf16* ws = reinterpret_cast<bf16*>(node.base(rank));
    uint32_t* flags = reinterpret_cast<uint32_t*>(node.base(rank) + ws_bytes);
    uint32_t state[2] = {0, 0};  // (phase, done): local to the rank
    std::vector<bf16> C(size_t(world) * max_rows * N);
    const int pc = world == 1 ? 0 : push_ctas;  // as the launch wrapper does
    for (int call = 0; call < calls; ++call) {
      const int rows = rows_of(call);
      const bf16* a = A[size_t(call) * world + rank].data();
      const bf16* b = B[rank].data();
      bf16* c = C.data();
      uint32_t* st = state;
      sim::launch(rank, ctas, 256, [=] {
        tdist::tut::tut_ag_gemm_kernel(ctx, a, b, c, ws, ws_bytes / 2 / sizeof(bf16), flags, blocks, rows, N, K, pc, tdist::PhaseState{st, st + 1});
      });
      for (int s = 0; s < world; ++s) {
        const bf16* as = A[size_t(call) * world + s].data();
        for (int i = 0; i < rows; ++i)
          for (int n = 0; n < N; ++n) {
            float acc = 0.f;
            for (int k = 0; k < K; ++k) acc = fmaf(__bfloat162float(as[size_t(i) * K + k]), __bfloat162float(b[size_t(n) * K + k]), acc);
            const float got = __bfloat162float(C[(size_t(s) * rows + i) * N + n]), want = __bfloat162float(__float2bfloat16(acc));
            if (fabsf(got - want) > 1e-2f * fmaxf(1.f, fabsf(want))) mismatch("ag_gemm", rank, call, (long(s) * rows + i) * N + n, got, want);
          }
      }
    }
  });
  printf("protocol_sim ag_gemm: %d ranks x %d calls, rows/rank %d%s N %d K %d, %d push + %d compute CTAs: %s\n", world, calls, max_rows,
         alternate == 1 ? " (odd calls: smaller)" : alternate == 2 ? " (even calls: smaller)" : "", N, K,
         world == 1 ? 0 : push_ctas, ctas - (world == 1 ? 0 : push_ctas), g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

// ------------------------------------------------------------------ tutorial 8: GEMM + ReduceScatter
// Rank r holds a K slice: partial_r = A_r[M, Kl] * B_r[N, Kl]^T; it must end up with the sum over ranks of rows
// [r * M/W, (r+1) * M/W) of the partials (each rounded to bf16 before the sum, as the kernel stores them).
int run_gemm_rs(int world, int max_M, int N, int Kl, int ctas, int calls, int alternate, int alt_div) {
  const int blocks = (max_M / world + 63) / 64, tiles_n = (N + 63) / 64;
  const size_t part_bytes = size_t(2) * max_M * N * sizeof(bf16);
  auto M_of = [=](int call) { return alternate && (call & 1) == (alternate & 1) ? max_M / alt_div : max_M; };
  const size_t flag_bytes = size_t(world) * blocks * tiles_n * sizeof(uint32_t);
  sim::Node node(world, part_bytes + flag_bytes + 64);
  std::vector<std::vector<bf16>> A(size_t(calls) * world), B(world);
  for (int r = 0; r < world; ++r) B[r] = random_bf16(size_t(N) * Kl, 555 + r, 0.25f);
  for (int c = 0; c < calls; ++c)
    for (int r = 0; r < world; ++r) A[size_t(c) * world + r] = random_bf16(size_t(M_of(c)) * Kl, 31 * c + r + 7, 1.f);
  sim::run_ranks(world, [&](int rank) {
    const tdist::SymmCtx ctx = node.ctx(rank);
    bf16* part = reinterpret_cast<bf16*>(node.base(rank));
    uint32_t* flags = reinterpret_cast<uint32_t*>(node.base(rank) + part_bytes);
    uint32_t state[2] = {0, 0};
    std::vector<bf16> out(size_t(max_M / world) * N);
    for (int call = 0; call < calls; ++call) {
      const int M = M_of(call), rows = M / world;
      const bf16* a = A[size_t(call) * world + rank].data();
      const bf16* b = B[rank].data();
      bf16* o = out.data();
      uint32_t* st = state;
      sim::launch(rank, ctas, 256, [=] {
        tdist::tut::tut_gemm_rs_kernel(ctx, a, b, o, part, part_bytes / 2 / sizeof(bf16), flags, blocks, M, N, Kl, tdist::PhaseState{st, st + 1});
      });
      for (int i = 0; i < rows; ++i)
        for (int n = 0; n < N; ++n) {
          float sum = 0.f;
          for (int p = 0; p < world; ++p) {
            const bf16* ap = A[size_t(call) * world + p].data() + (size_t(rank) * rows + i) * Kl;
            const bf16* bp = B[p].data() + size_t(n) * Kl;
            float acc = 0.f;
            for (int k = 0; k < Kl; ++k) acc = fmaf(__bfloat162float(ap[k]), __bfloat162float(bp[k]), acc);
            sum += __bfloat162float(__float2bfloat16(acc));
          }
          const float got = __bfloat162float(out[size_t(i) * N + n]);
          if (fabsf(got - sum) > 2e-2f * fmaxf(1.f, fabsf(sum))) mismatch("gemm_rs", rank, call, long(i) * N + n, got, sum);
        }
    }
  });
  printf("protocol_sim gemm_rs: %d ranks x %d calls, M %d%s N %d K/W %d, %d CTAs: %s\n", world, calls, max_M,
         alternate == 1 ? " (odd calls: smaller)" : alternate == 2 ? " (even calls: smaller)" : "", N, Kl, ctas,
         g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

}  // namespace

int main(int argc, char** argv) {
  auto arg = [&](int i, int dflt) { return argc > i ? atoi(argv[i]) : dflt; };
  const std::string what = argc > 1 ? argv[1] : "queue";
  if (what == "queue") return run_queue(arg(2, 3), arg(3, 11), arg(4, 300), arg(5, 3), arg(6, 3));
  if (what == "ag_gemm") return run_ag_gemm(arg(2, 3), arg(3, 96), arg(4, 72), arg(5, 32), arg(6, 2), arg(7, 5), arg(8, 3), arg(9, 0), arg(10, 2));
  if (what == "gemm_rs") return run_gemm_rs(arg(2, 3), arg(3, 3 * 96), arg(4, 72), arg(5, 32), arg(6, 4), arg(7, 3), arg(8, 0), arg(9, 2));
  fprintf(stderr, "usage: protocol_sim queue|ag_gemm|gemm_rs [sizes...]\n");
  return 2;
}
