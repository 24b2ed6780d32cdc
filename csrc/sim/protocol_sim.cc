// protocol_sim.cc — runs the distributed kernels of this library on the CPU, one OS thread per CTA (its CUDA threads are
// fibers) and one "GPU" per rank inside ONE process, under ThreadSanitizer. (SURVEY.md section 5 asks for a flag-protocol
// checker; the reference has none — its substitute is noise injection on real GPUs.) docs/simulator.md is the guide.
//
// The kernel sources are compiled UNMODIFIED up to three mechanical steps done by tests/test_cpu_protocol_sim.py when it
// stages them: the files are cut before their launch wrappers (which use <<< >>> or the runtime API), `__shared__ T name[..];`
// becomes a reference into a per-CTA buffer, and the two inline-asm functions of the megakernel call their emulations.
// csrc/sim/shim/device/ptx.cuh stands in for csrc/device/ptx.cuh and csrc/sim/stub/*.h for the CUDA headers: flags become C++
// atomics with the memory orders the device code uses, payload accesses stay plain, __syncthreads() yields between fibers,
// bulk / tiled TMA copies are performed at issue. ThreadSanitizer then reports any payload access that is not ordered by the
// flag protocol, a lost wake-up shows up as the wait bound, and the results are compared with a host computation. Ranks issue
// their calls independently (no barrier between calls), as streams on different GPUs do. TDIST_SIM_SLOW_RANK=r
// (+ TDIST_SIM_SLOW_MS) delays every flag write of rank r: the skew that opens the windows a protocol must survive.
//
//   protocol_sim queue   <ranks> <items> <item_vecs> <slots> <calls>
//   protocol_sim ag_gemm <ranks> <rows> <N> <K> <push_ctas> <ctas> <calls> [alternate: 1 = odd calls use rows / div, 2 = even calls] [div = 2]
//   protocol_sim coll <op> <ranks> <vecs> <ctas> <calls> [alternate]   op: ar1 ar2 ar1mc ar2mc ar1mc_inplace ag_pull ag_push
//                                                                    ag_mc ag_ll ag_llmc rs rsmc barrier selftest
//   protocol_sim moe_utils <tokens> <topk> <experts> <ranks> <row_vecs>      (one simulated GPU: functional check)
//   protocol_sim moe_rs <ranks> <tokens> <topk> <row_vecs> <ctas> <calls> [alternate]
//   protocol_sim a2a <ranks> <max rows per peer> <row_vecs> <experts per rank> <ctas per peer> <calls>
//   protocol_sim ep <ranks> <tokens> <hidden> <experts> <ctas> <calls> [fp8]
//   protocol_sim fused rs|rsl|rs2|ar1|ar2 <ranks> <M> <N> <ctas> <calls> [in-switch reduction 0/1] [alternate] [fixed flag rows 0/1]   (production GEMM comm side)
//   protocol_sim decode <batch> <Hq> <Hkv> <max cached tokens> <kv splits> <page size, 0 = dense> <hidden> <intermediate> <vocab> [fp16]
//                                                                    (one simulated GPU: one decode step, every kernel against a host oracle)
//   protocol_sim mega <ranks> <CTAs> <layers> <batch> <steps> <graph: 0 fused, 1 op by op, 2 fused + one-task attention> <in-switch all-reduce 0/1>
//                     <KV heads per rank> <query heads per KV head> <MLP columns per rank> <KV splits>     (the decode megakernel + its native builder)
//   protocol_sim fused ag|ag2 <ranks> <rows per rank> <N> <K> <push CTAs> <CTAs> <calls> [alternate] [row slices, 0 = point-to-point ring] [K slabs] [fixed flag rows 0/1]
//                                                                    (production AG + GEMM; ag2: the CTA-pair kernel)
//   protocol_sim gemm_rs <ranks> <M> <N> <Kl> <ctas> <calls> [alternate: 1 = odd calls use M / div, 2 = even calls] [div = 2]
#include "sim_runtime.h"

#include "tutorials/notify_wait_tutorial.cu"  // tdist::tut::tut_queue_kernel
#include "tutorials/overlap_tutorial.cu"      // tdist::tut::tut_ag_gemm_kernel, tut_gemm_rs_kernel
#include "kernels/collectives.cu"             // PRODUCTION: barrier, all-reduce x4, all-gather x5, reduce-scatter x2
#include "kernels/primitives_selftest.cu"     // PRODUCTION: every NVSHMEM-style device primitive of symm.cuh in one kernel
#include "kernels/moe.cu"                     // PRODUCTION: MoE align / sort, tile swizzle, gathers, top-k reduce (+ reduce-scatter)
#include "kernels/ep_a2a.cu"                  // PRODUCTION: low-latency all-to-all, EP counts / dispatch (bf16, fp8) / combine
#include "kernels/decode.cu"                  // PRODUCTION: decode path (GEMV, norms, rotary + KV append, split-KV attention, sampling)
#include "kernels/gemm_sm100.cuh"             // PRODUCTION: parameter block, tile order and communication bodies of the fused GEMM
#include "kernels/gemm_sm100_2cta.cuh"        // PRODUCTION: tile order of the CTA-pair kernel
#include "mega/megakernel.cu"                 // PRODUCTION: the decode megakernel (interpreter + every task body)

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


// ------------------------------------------------------------------ production collectives (csrc/kernels/collectives.cu)
// fp32 payload with small integer values: sums are exact in any order, so results are compared bit for bit. `vecs` =
// 16-byte vectors of the message (all-reduce), of one rank's shard (all-gather) or of one segment (reduce-scatter); with
// `alternate` odd (1) or even (2) calls use a third of it on the SAME workspace, flags and phase counters.
int run_collective(const std::string& op, int world, int max_vecs, int ctas, int calls, int alternate) {
  using tdist::F32;
  const bool mc = op.find("mc") != std::string::npos;
  const bool is_ag = op.rfind("ag_", 0) == 0, is_rs = op.rfind("rs", 0) == 0, is_ll = op.find("ll") != std::string::npos;
  const size_t W = world;
  // symmetric layout per rank: [workspace | flags]
  size_t ws_vecs;
  if (is_ag) ws_vecs = is_ll ? 2 * W * size_t(max_vecs) * 2 : 2 * W * size_t(max_vecs);  // LL: packets, two parity halves
  else if (is_rs) ws_vecs = 2 * W * size_t(max_vecs);
  else ws_vecs = 2 * (W * (size_t(max_vecs) / W + 1 + 2 * ctas) + size_t(max_vecs)) + 2 * W * size_t(max_vecs);
  const size_t ws_bytes = ws_vecs * 16, flag_words = size_t(ctas + 1) * TDIST_MAX_RANKS;
  sim::Node node(world, ws_bytes + flag_words * 4 + 64, mc);
  if (mc && world == 1) {
    printf("protocol_sim coll %s: needs more than one rank\n", op.c_str());
    return 0;
  }
  auto vecs_of = [=](int call) { return alternate && (call & 1) == (alternate & 1) ? std::max(max_vecs / 3, 1) : max_vecs; };
  auto val = [](int call, int rank, size_t i) { return float((call * 7 + rank * 3 + int(i % 11)) % 32); };
  sim::run_ranks(world, [&](int rank) {
    const tdist::SymmCtx ctx = node.ctx(rank);
    int4* ws = reinterpret_cast<int4*>(node.base(rank));
    uint32_t* flags = reinterpret_cast<uint32_t*>(node.base(rank) + ws_bytes);
    uint32_t state[2] = {0, 0};
    const tdist::PhaseState ps{state, state + 1};
    for (int call = 0; call < calls; ++call) {
      const size_t n = vecs_of(call);
      const int grid = int(std::min<size_t>(ctas, std::max<size_t>(1, n / 8)));  // as pick_grid: fewer CTAs for small messages
      if (op == "barrier") {
        sim::launch(rank, 1, 32, [=] { tdist::barrier_all_kernel(ctx, flags, ps); });
        continue;
      }
      if (is_ag) {
        std::vector<float> in(n * 4);
        for (size_t i = 0; i < n * 4; ++i) in[i] = val(call, rank, i);
        const float* got;
        std::vector<float> out;
        if (is_ll) {  // packets go through the context's buffer (two parity halves inside the kernel), result is local
          out.assign(W * n * 4, -1.f);
          const int4* inp = reinterpret_cast<const int4*>(in.data());
          int4* outp = reinterpret_cast<int4*>(out.data());
          const size_t n64 = n * 2, parity_pkts = ws_bytes / 16 / 2;
          const int use_mc = op == "ag_llmc";
          sim::launch(rank, grid, 64, [=] {
            tdist::allgather_ll_kernel(ctx, reinterpret_cast<const uint2*>(inp), ws, reinterpret_cast<uint2*>(outp), n64, parity_pkts, ps, use_mc);
          });
          got = out.data();
        } else {
          // in place, as fast_allgather uses it: slot `rank` of a symmetric [W][n] buffer is filled by the caller, all slots
          // are valid on return. The BUFFER is the caller's: consecutive calls here use two buffers alternately (a caller
          // that reuses one buffer puts a barrier between the calls, as ag_group_gemm does).
          int4* symm = ws + size_t(call & 1) * W * max_vecs;
          int4* mine = symm + size_t(rank) * n;
          memcpy(mine, in.data(), n * 16);
          const int mode = op == "ag_pull" ? 0 : op == "ag_push" ? 1 : 2;
          sim::launch(rank, grid, 64, [=] { tdist::allgather_kernel(ctx, mine, symm, static_cast<int4*>(nullptr), n, flags, ps, mode); });
          got = reinterpret_cast<const float*>(symm);
        }
        for (size_t s = 0; s < W; ++s)
          for (size_t i = 0; i < n * 4; ++i)
            if (got[s * n * 4 + i] != val(call, int(s), i)) mismatch(op.c_str(), rank, call, long(s * n * 4 + i), got[s * n * 4 + i], val(call, int(s), i));
        continue;
      }
      if (is_rs) {
        // in: [W][n] (segment s goes to rank s); out: [n] = sum over ranks of their segment `rank`
        std::vector<float> in(W * n * 4), out(n * 4, -1.f);
        for (size_t sgm = 0; sgm < W; ++sgm)
          for (size_t i = 0; i < n * 4; ++i) in[sgm * n * 4 + i] = val(call, rank, sgm * 5 + i);
        const int4* inp = reinterpret_cast<const int4*>(in.data());
        int4* outp = reinterpret_cast<int4*>(out.data());
        const size_t parity_vecs = ws_bytes / 16 / 2;
        if (op == "rsmc")
          sim::launch(rank, grid, 64, [=] { tdist::reduce_scatter_mc_kernel<F32>(ctx, inp, ws, outp, n, parity_vecs, flags, ps); });
        else
          sim::launch(rank, grid, 64, [=] { tdist::reduce_scatter_kernel<F32>(ctx, inp, ws, outp, n, parity_vecs, flags, ps); });
        for (size_t i = 0; i < n * 4; ++i) {
          float want = 0.f;
          for (size_t r = 0; r < W; ++r) want += val(call, int(r), size_t(rank) * 5 + i);
          if (out[i] != want) mismatch(op.c_str(), rank, call, long(i), out[i], want);
        }
        continue;
      }
      if (op == "selftest") {
        // csrc/kernels/primitives_selftest.cu: buf [3W + 2] int4, flags [2W + 1]; the kernel writes `epoch` into buf[3W].x
        // when every primitive delivered what it should
        int4* buf = ws;
        const uint32_t epoch = call + 1;
        sim::launch(rank, 1, 128, [=] { tdist::primitives_selftest_kernel(ctx, buf, flags, epoch); });
        const int res = buf[3 * W].x;
        if (res != int(epoch)) mismatch("selftest", rank, call, 0, float(res), float(epoch));
        // the test script puts a barrier between epochs (nvshmem_barrier_all_on_stream): buffers are reused
        sim::launch(rank, 1, 32, [=] { tdist::barrier_all_kernel(ctx, flags + 2 * W + 8, ps); });
        continue;
      }
      // all-reduce: in [n], out [n] = sum over ranks
      std::vector<float> in(n * 4), out(n * 4, -1.f);
      for (size_t i = 0; i < n * 4; ++i) in[i] = val(call, rank, i);
      const int4* inp = reinterpret_cast<const int4*>(in.data());
      int4* outp = reinterpret_cast<int4*>(out.data());
      if (op == "ar1") {
        const size_t slot_vecs = ws_bytes / 16 / 2 / W;
        sim::launch(rank, grid, 64, [=] { tdist::ar_one_shot_push_kernel<F32>(ctx, inp, outp, ws, slot_vecs, flags, ps, n); });
      } else if (op == "ar2") {
        const size_t shard_cap = (n + W - 1) / W + 2 * size_t(grid), parity_vecs = ws_bytes / 16 / 2;
        sim::launch(rank, grid, 64, [=] { tdist::ar_two_shot_push_kernel<F32>(ctx, inp, outp, ws, parity_vecs, shard_cap, flags, ps, n); });
      } else if (op == "ar1mc" || op == "ar2mc") {
        const size_t parity_vecs = ws_bytes / 16 / 2;
        if (op == "ar1mc")
          sim::launch(rank, grid, 64, [=] { tdist::ar_one_shot_mc_kernel<F32>(ctx, inp, outp, ws, parity_vecs, flags, ps, n, 0); });
        else
          sim::launch(rank, grid, 64, [=] { tdist::ar_two_shot_mc_kernel<F32>(ctx, inp, outp, ws, parity_vecs, flags, ps, n, 0); });
      } else if (op == "ar1mc_inplace" || op == "ar2mc_inplace") {
        // the caller's data already lives in the symmetric buffer (GEMV + all-reduce of the decode path): no copy-in, a
        // closing barrier keeps peers from overwriting it while others still read
        memcpy(ws, in.data(), n * 16);
        if (op == "ar1mc_inplace")
          sim::launch(rank, grid, 64, [=] { tdist::ar_one_shot_mc_kernel<F32>(ctx, ws, outp, ws, 0, flags, ps, n, 1); });
        else
          sim::launch(rank, grid, 64, [=] { tdist::ar_two_shot_mc_kernel<F32>(ctx, ws, outp, ws, 0, flags, ps, n, 1); });
      } else {
        fprintf(stderr, "unknown collective %s\n", op.c_str());
        exit(2);
      }
      for (size_t i = 0; i < n * 4; ++i) {
        float want = 0.f;
        for (size_t r = 0; r < W; ++r) want += val(call, int(r), i);
        if (out[i] != want) mismatch(op.c_str(), rank, call, long(i), out[i], want);
      }
    }
  });
  printf("protocol_sim coll %s: %d ranks x %d calls, %d vecs%s, up to %d CTAs: %s\n", op.c_str(), world, calls, max_vecs,
         alternate ? " (alternating with a third)" : "", ctas, g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

}  // namespace

#include "moe_ep_scenarios.h"
#include "fused_gemm_scenarios.h"
#include "decode_scenarios.h"
#include "mega_scenarios.h"

int main(int argc, char** argv) {
  auto arg = [&](int i, int dflt) { return argc > i ? atoi(argv[i]) : dflt; };
  const std::string what = argc > 1 ? argv[1] : "queue";
  if (what == "queue") return run_queue(arg(2, 3), arg(3, 11), arg(4, 300), arg(5, 3), arg(6, 3));
  if (what == "ag_gemm") return run_ag_gemm(arg(2, 3), arg(3, 96), arg(4, 72), arg(5, 32), arg(6, 2), arg(7, 5), arg(8, 3), arg(9, 0), arg(10, 2));
  if (what == "coll") return run_collective(argc > 2 ? argv[2] : "ar1", arg(3, 3), arg(4, 96), arg(5, 3), arg(6, 3), arg(7, 0));
  if (what == "moe_utils") return run_moe_utils(arg(2, 96), arg(3, 2), arg(4, 5), arg(5, 3), arg(6, 2));
  if (what == "moe_rs") return run_moe_rs(arg(2, 3), arg(3, 48), arg(4, 2), arg(5, 3), arg(6, 3), arg(7, 4), arg(8, 0));
  if (what == "a2a") return run_a2a(arg(2, 3), arg(3, 24), arg(4, 3), arg(5, 2), arg(6, 2), arg(7, 4));
  if (what == "ep") return run_ep(arg(2, 3), arg(3, 12), arg(4, 256), arg(5, 6), arg(6, 2), arg(7, 3), arg(8, 0));
  if (what == "fused" && argc > 2 && (std::string(argv[2]) == "ag" || std::string(argv[2]) == "ag2"))
    return run_fused_ag(arg(3, 3), arg(4, 192), arg(5, 128), arg(6, 192), arg(7, 2), arg(8, 6), arg(9, 3), arg(10, 0), arg(11, 2), arg(12, 2), arg(13, 0),
                        std::string(argv[2]) == "ag2");
  if (what == "fused") return run_fused(argc > 2 ? argv[2] : "rs", arg(3, 3), arg(4, 3 * 256), arg(5, 128), arg(6, 3), arg(7, 3), arg(8, 1), arg(9, 0), arg(10, 0));
  if (what == "decode") return run_decode(arg(2, 3), arg(3, 14), arg(4, 2), arg(5, 70), arg(6, 5), arg(7, 0), arg(8, 128), arg(9, 64), arg(10, 300), arg(11, 0));
  if (what == "mega") return run_mega(arg(2, 2), arg(3, 4), arg(4, 2), arg(5, 2), arg(6, 3), arg(7, 0), arg(8, 1), arg(9, 1), arg(10, 2), arg(11, 128), arg(12, 3));
  if (what == "gemm_rs") return run_gemm_rs(arg(2, 3), arg(3, 3 * 96), arg(4, 72), arg(5, 32), arg(6, 4), arg(7, 3), arg(8, 0), arg(9, 2));
  fprintf(stderr, "usage: protocol_sim queue|ag_gemm|gemm_rs|coll|moe_utils|moe_rs|a2a|ep|fused|decode|mega [sizes...]\n");
  return 2;
}

// the native task-graph builder and scheduler of the megakernel, as it is (host code; its "device" memory is host memory here
// and its launch goes to tdist_mega_launch in mega_scenarios.h). Last: it ends with file-scope using-directives.
#include "mega/builder.cc"
