// moe_ep_scenarios.h — simulator scenarios for the PRODUCTION MoE / expert-parallel kernels (csrc/kernels/moe.cu,
// ep_a2a.cu). Included by protocol_sim.cc after the kernel sources.
#pragma once

namespace {

unsigned lcg(unsigned& s) { return s = s * 1664525u + 1013904223u; }

// ------------------------------------------------------------------ MoE index kernels on one simulated GPU
// moe_align_sort -> moe_tile_swizzle -> moe_gather_rows (+ the stage-filtered form) -> moe_topk_reduce, each against a host
// restatement. No communication: this is a functional check of kernels that otherwise only run on a GPU.
int run_moe_utils(int tokens, int topk, int experts, int world, int row_vecs) {
  const int block_m = 128, pairs = tokens * topk;
  const int max_rows = (pairs + std::min(experts, pairs) * (block_m - 1) + block_m - 1) / block_m * block_m;
  unsigned seed = 99;
  std::vector<int> ids(pairs);
  for (int t = 0; t < tokens; ++t)
    for (int k = 0; k < topk; ++k) ids[t * topk + k] = (lcg(seed) >> 8) % experts;
  std::vector<int> sorted(max_rows, -7), pair_to_row(pairs, -7), block_expert(max_rows / block_m, -7), meta(2, -7);
  sim::launch(0, 1, 256, [&] {
    tdist::moe_align_sort_kernel(ids.data(), pairs, experts, block_m, sorted.data(), pair_to_row.data(), block_expert.data(), meta.data(), max_rows);
  });
  // invariants of the layout
  std::vector<int> cnt(experts, 0);
  for (int p : ids) ++cnt[p];
  int rows = 0;
  for (int e = 0; e < experts; ++e) rows += (cnt[e] + block_m - 1) / block_m * block_m;
  if (meta[0] != rows || meta[1] != rows / block_m) mismatch("moe_align_sort meta", 0, 0, 0, float(meta[0]), float(rows));
  std::vector<int> seen(pairs, 0);
  for (int i = 0; i < max_rows; ++i) {
    const int p = sorted[i], be = block_expert[i / block_m];
    if (i >= rows) {
      if (p != pairs || be != -1) mismatch("moe_align_sort tail", 0, 0, i, float(p), float(pairs));
      continue;
    }
    if (p == pairs) continue;  // padding row of a block
    if (p < 0 || p > pairs || ids[p] != be || pair_to_row[p] != i) mismatch("moe_align_sort row", 0, 0, i, float(p), float(be));
    else ++seen[p];
  }
  for (int p = 0; p < pairs; ++p)
    if (seen[p] != 1) mismatch("moe_align_sort coverage", 0, 0, p, float(seen[p]), 1.f);
  // tile swizzle as rank `r` of `world`
  const int tokens_per_rank = tokens / world, nblk = max_rows / block_m;
  for (int r = 0; r < world; ++r) {
    std::vector<int> order(nblk, -1), s0(nblk, -1), s1(nblk, -1), stage_count(TDIST_MAX_RANKS, 0);
    sim::launch(0, 1, 256, [&] {
      moe_tile_swizzle_kernel(sorted.data(), meta.data(), pairs, topk, tokens_per_rank, r, world, block_m, order.data(), s0.data(), s1.data(),
                              stage_count.data(), nblk);
    });
    std::vector<int> want_order;
    for (int st = 0; st < world; ++st)
      for (int b = 0; b < meta[1]; ++b) {
        int lo = world, hi = -1;
        for (int i = 0; i < block_m; ++i) {
          const int p = sorted[b * block_m + i];
          if (p < pairs) {
            const int s = ((p / topk) / tokens_per_rank - r + world) % world;
            lo = std::min(lo, s);
            hi = std::max(hi, s);
          }
        }
        if (hi < 0) lo = hi = 0;
        if (st == 0 && (s0[b] != lo || s1[b] != hi)) mismatch("moe_tile_swizzle range", r, 0, b, float(s1[b]), float(hi));
        if (hi == st) want_order.push_back(b);
      }
    for (size_t i = 0; i < want_order.size(); ++i)
      if (order[i] != want_order[i]) mismatch("moe_tile_swizzle order", r, 0, long(i), float(order[i]), float(want_order[i]));
    // stage-filtered gather: the stages together fill exactly what the one-shot gather fills
    std::vector<float> x(size_t(tokens) * row_vecs * 4), full(size_t(max_rows) * row_vecs * 4, -5.f), staged(full.size(), -5.f);
    for (size_t i = 0; i < x.size(); ++i) x[i] = float(i % 251);
    const int4* xs = reinterpret_cast<const int4*>(x.data());
    int4* fd = reinterpret_cast<int4*>(full.data());
    int4* sd = reinterpret_cast<int4*>(staged.data());
    sim::launch(0, 3, 256, [&] { tdist::moe_gather_rows_kernel(xs, fd, sorted.data(), meta.data(), pairs, topk, row_vecs); });
    for (int st = 0; st < world; ++st)
      sim::launch(0, 3, 256, [&] { tdist::moe_gather_rows_stage_kernel(xs, sd, sorted.data(), meta.data(), s1.data(), st, block_m, pairs, topk, row_vecs); });
    for (size_t i = 0; i < size_t(meta[0]) * row_vecs * 4; ++i) {
      const int row = int(i / (row_vecs * 4)), p = sorted[row];
      const float want = p < pairs ? x[size_t(p / topk) * row_vecs * 4 + i % (row_vecs * 4)] : 0.f;
      if (full[i] != want) mismatch("moe_gather_rows", r, 0, long(i), full[i], want);
      if (staged[i] != want) mismatch("moe_gather_rows_stage", r, 0, long(i), staged[i], want);
    }
  }
  // top-k weighted reduce (fp32: exact with small integers)
  std::vector<float> y(size_t(max_rows) * row_vecs * 4), w(pairs), out(size_t(tokens) * row_vecs * 4, -1.f);
  for (size_t i = 0; i < y.size(); ++i) y[i] = float(i % 13);
  for (int p = 0; p < pairs; ++p) w[p] = float(1 + p % 3);
  sim::launch(0, 2, 256, [&] {
    tdist::moe_topk_reduce_kernel<tdist::F32>(reinterpret_cast<const int4*>(y.data()), reinterpret_cast<int4*>(out.data()), pair_to_row.data(),
                                              w.data(), tokens, topk, row_vecs);
  });
  for (int t = 0; t < tokens; ++t)
    for (int j = 0; j < row_vecs * 4; ++j) {
      float want = 0.f;
      for (int k = 0; k < topk; ++k) want += w[t * topk + k] * y[size_t(pair_to_row[t * topk + k]) * row_vecs * 4 + j];
      if (out[size_t(t) * row_vecs * 4 + j] != want) mismatch("moe_topk_reduce", 0, 0, long(t) * row_vecs * 4 + j, out[size_t(t) * row_vecs * 4 + j], want);
    }
  printf("protocol_sim moe_utils: %d tokens top-%d over %d experts, layout as %d ranks: %s\n", tokens, topk, experts, world,
         g_errors ? "MISMATCH" : "ok");
  return g_errors ? 1 : 0;
}

// ------------------------------------------------------------------ fused top-k reduce + reduce-scatter (moe_reduce_rs_kernel)
// Every rank holds partial outputs y (rows in an arbitrary "sorted" order) for ALL tokens; token segment s belongs to rank s.
// Calls alternate between `tokens` and tokens / 2 on the same scatter workspace when `alternate` is set.
int run_moe_rs(int world, int max_tokens, int topk, int row_vecs, int ctas, int calls, int alternate) {
  const size_t W = world;
  const size_t ws_bytes = 2 * size_t(max_tokens) * row_vecs * 16;  // [2 parity][W][tokens / W][row]
  sim::Node node(world, ws_bytes + size_t(ctas + 1) * TDIST_MAX_RANKS * 4 + 64);
  auto tokens_of = [=](int call) { return alternate && (call & 1) == (alternate & 1) ? max_tokens / 2 : max_tokens; };
  auto yval = [](int call, int rank, size_t i) { return float((call * 5 + rank * 3 + int(i % 7)) % 16); };
  sim::run_ranks(world, [&](int rank) {
    const tdist::SymmCtx ctx = node.ctx(rank);
    int4* scat = reinterpret_cast<int4*>(node.base(rank));
    uint32_t* flags = reinterpret_cast<uint32_t*>(node.base(rank) + ws_bytes);
    uint32_t state[2] = {0, 0};
    const tdist::PhaseState ps{state, state + 1};
    for (int call = 0; call < calls; ++call) {
      const int tokens = tokens_of(call), pairs = tokens * topk, seg = tokens / world;
      std::vector<int> pair_to_row(pairs);
      for (int p = 0; p < pairs; ++p) pair_to_row[p] = (p * 7 + 3) % pairs;  // a permutation (7 and pairs coprime for our sizes)
      std::vector<float> y(size_t(pairs) * row_vecs * 4), w(pairs), out(size_t(seg) * row_vecs * 4, -1.f);
      for (size_t i = 0; i < y.size(); ++i) y[i] = yval(call, rank, i);
      for (int p = 0; p < pairs; ++p) w[p] = float(1 + p % 2);
      const int4* yp = reinterpret_cast<const int4*>(y.data());
      int4* op = reinterpret_cast<int4*>(out.data());
      const int* ptr = pair_to_row.data();
      const float* wp = w.data();
      const size_t parity_vecs = ws_bytes / 16 / 2;
      const int grid = int(std::min<size_t>(ctas, std::max<size_t>(1, size_t(seg) * row_vecs / 8)));
      sim::launch(rank, grid, 64, [=] {
        tdist::moe_reduce_rs_kernel<tdist::F32>(ctx, yp, scat, op, ptr, wp, tokens, topk, row_vecs, parity_vecs, flags, ps);
      });
      for (int t = 0; t < seg; ++t)
        for (int j = 0; j < row_vecs * 4; ++j) {
          float want = 0.f;
          const int tok = rank * seg + t;
          for (size_t r = 0; r < W; ++r)
            for (int k = 0; k < topk; ++k)
              want += w[tok * topk + k] * yval(call, int(r), size_t(pair_to_row[tok * topk + k]) * row_vecs * 4 + j);
          const float got = out[size_t(t) * row_vecs * 4 + j];
          if (got != want) mismatch("moe_reduce_rs", rank, call, long(t) * row_vecs * 4 + j, got, want);
        }
    }
  });
  printf("protocol_sim moe_rs: %d ranks x %d calls, %d tokens%s top-%d, %d vecs per row: %s\n", world, calls, max_tokens,
         alternate ? " (alternating with half)" : "", topk, row_vecs, g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

// ------------------------------------------------------------------ low-latency all-to-all (a2a_kernel)
// Rank r sends rows [cumsum[p * epr], cumsum[(p + 1) * epr]) to peer p, with per-expert splits and a per-row scale.
int run_a2a(int world, int max_m, int row_vecs, int epr, int ctas_per_peer, int calls) {
  const size_t W = world, row_bytes = size_t(row_vecs) * 16;
  const size_t data_bytes = 2 * W * max_m * row_bytes, split_bytes = 2 * W * epr * 4, scale_bytes = 2 * W * max_m * 4;
  sim::Node node(world, data_bytes + split_bytes + scale_bytes + 2 * TDIST_MAX_RANKS * 4 + 64);
  auto rows_to = [=](int call, int src, int dst, int e) { return (call + src * 2 + dst + e) % (max_m / std::max(epr, 1) + 1); };
  sim::run_ranks(world, [&](int rank) {
    const tdist::SymmCtx ctx = node.ctx(rank);
    char* recv_data = node.base(rank);
    int* recv_splits = reinterpret_cast<int*>(recv_data + data_bytes);
    float* recv_scale = reinterpret_cast<float*>(recv_data + data_bytes + split_bytes);
    uint32_t* flags = reinterpret_cast<uint32_t*>(recv_data + data_bytes + split_bytes + scale_bytes);
    uint32_t state[2] = {0, 0};
    const tdist::PhaseState ps{state, state + 1};
    for (int call = 0; call < calls; ++call) {
      std::vector<int> cumsum(W * epr + 1, 0);
      for (size_t p = 0; p < W; ++p)
        for (int e = 0; e < epr; ++e) cumsum[p * epr + e + 1] = cumsum[p * epr + e] + rows_to(call, rank, int(p), e);
      const int total = cumsum[W * epr];
      std::vector<float> send(size_t(total) * row_vecs * 4), scale(std::max(total, 1));
      for (int i = 0; i < total; ++i) {
        scale[i] = float(rank * 1000 + i);
        for (int j = 0; j < row_vecs * 4; ++j) send[size_t(i) * row_vecs * 4 + j] = float(call * 100000 + rank * 10000 + i * 8 + j % 8);
      }
      const char* sp = reinterpret_cast<const char*>(send.data());
      const int* cs = cumsum.data();
      const float* sc = scale.data();
      sim::launch(rank, world * ctas_per_peer, 64, [=] {
        tdist::a2a_kernel(ctx, sp, cs, sc, recv_data, recv_splits, recv_scale, flags, ps, max_m, int(row_bytes), epr, ctas_per_peer);
      });
      const int parity = (call + 1) & 1;  // the kernel numbers its calls from 1
      for (size_t s = 0; s < W; ++s) {
        int off_src = 0;  // where the rows for me start in s's send buffer
        for (int p = 0; p < rank; ++p)
          for (int e = 0; e < epr; ++e) off_src += rows_to(call, int(s), p, e);
        int n = 0;
        for (int e = 0; e < epr; ++e) {
          const int want = rows_to(call, int(s), rank, e), got = recv_splits[parity * W * epr + s * epr + e];
          if (got != want) mismatch("a2a splits", rank, call, long(s * epr + e), float(got), float(want));
          n += want;
        }
        const float* rows = reinterpret_cast<const float*>(recv_data + (size_t(parity) * W + s) * max_m * row_bytes);
        const float* scl = recv_scale + (size_t(parity) * W + s) * max_m;
        for (int i = 0; i < n; ++i) {
          if (scl[i] != float(s * 1000 + off_src + i)) mismatch("a2a scale", rank, call, i, scl[i], float(s * 1000 + off_src + i));
          for (int j = 0; j < row_vecs * 4; ++j) {
            const float want = float(call * 100000 + int(s) * 10000 + (off_src + i) * 8 + j % 8);
            if (rows[size_t(i) * row_vecs * 4 + j] != want) mismatch("a2a data", rank, call, long(i) * row_vecs * 4 + j, rows[size_t(i) * row_vecs * 4 + j], want);
          }
        }
      }
    }
  });
  printf("protocol_sim a2a: %d ranks x %d calls, up to %d rows per peer, %d experts per rank, %d CTAs per peer: %s\n", world, calls,
         max_m, epr, ctas_per_peer, g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

// ------------------------------------------------------------------ EP dispatch / combine (ep_counts, ep_dispatch, ep_combine)
// One MoE layer step per call, as EPAll2AllLayer runs it: counts exchange -> dispatch (optionally fp8) -> "experts" (identity:
// the received rows are copied into the symmetric expert-output buffer) -> combine with weights 1. Without fp8 the result
// must be exactly topk * x (bf16 doubling is exact for topk = 2); the receive layout and the provenance of every row are
// checked as well. Buffers, flag rows and the shared phase counter are laid out as in triton_dist/kernels/nvidia/ep_a2a.py.
int run_ep(int world, int tokens, int hidden, int experts, int ctas, int calls, int fp8) {
  using bf16 = __nv_bfloat16;
  const int topk = 2, pairs = tokens * topk, epr = experts / world, cap = pairs * world;
  const size_t W = world;
  const size_t off_counts = 0, off_recv = off_counts + W * experts * 4, off_scale = off_recv + size_t(cap) * hidden * 2,
               off_src = off_scale + size_t(cap) * (hidden / 128) * 4, off_out = off_src + size_t(cap) * 4,
               off_flags = off_out + size_t(cap) * hidden * 2, total = off_flags + 4 * TDIST_MAX_RANKS * 4;
  sim::Node node(world, total + 64);
  auto xval = [](int call, int rank, int t, int j) { return float((call * 3 + rank * 5 + t * 7 + j) % 64); };  // exact in bf16
  auto idval = [=](int call, int rank, int t, int k) { return (call + rank * 3 + t * 5 + k * (experts / 2 + 1)) % experts; };
  sim::run_ranks(world, [&](int rank) {
    const tdist::SymmCtx ctx = node.ctx(rank);
    char* base = node.base(rank);
    int* full_counts = reinterpret_cast<int*>(base + off_counts);
    char* recv_buf = base + off_recv;
    float* recv_scale = reinterpret_cast<float*>(base + off_scale);
    int* recv_src = reinterpret_cast<int*>(base + off_src);
    bf16* expert_out = reinterpret_cast<bf16*>(base + off_out);
    uint32_t* flags = reinterpret_cast<uint32_t*>(base + off_flags);
    uint32_t state[2] = {0, 0};
    uint32_t grid_counter[4] = {0, 0, 0, 0};
    std::vector<int> scatter_idx(pairs), cursors(experts, 0), recv_counts(epr, 0), num_recv(1, 0);
    const tdist::PhaseState ps{state, state + 1};
    for (int call = 0; call < calls; ++call) {
      std::vector<bf16> x(size_t(tokens) * hidden), out(size_t(tokens) * hidden);
      std::vector<int> ids(pairs);
      for (int t = 0; t < tokens; ++t) {
        for (int j = 0; j < hidden; ++j) x[size_t(t) * hidden + j] = __float2bfloat16(xval(call, rank, t, j));
        for (int k = 0; k < topk; ++k) ids[t * topk + k] = idval(call, rank, t, k);
      }
      const bf16* xp = x.data();
      const int* idp = ids.data();
      int *sip = scatter_idx.data(), *cup = cursors.data(), *rcp = recv_counts.data(), *nrp = num_recv.data();
      uint32_t* gc = grid_counter;
      sim::launch(rank, 1, 256, [=] { tdist::ep_counts_kernel(ctx, idp, pairs, experts, full_counts, flags, ps); });
      sim::launch(rank, ctas, 256, [=] {
        tdist::ep_dispatch_kernel<bf16>(ctx, xp, idp, full_counts, recv_buf, recv_scale, recv_src, sip, cup, rcp, nrp, tokens, topk, hidden,
                                        experts, fp8, flags + 16, gc, ps);
      });
      // receive layout: rows grouped by local expert; every row carries what its provenance says
      int expect_total = 0;
      for (int e = 0; e < epr; ++e) {
        int n = 0;
        for (size_t s = 0; s < W; ++s)
          for (int t = 0; t < tokens; ++t)
            for (int k = 0; k < topk; ++k) n += idval(call, int(s), t, k) == rank * epr + e;
        if (recv_counts[e] != n) mismatch("ep recv_expert_counts", rank, call, e, float(recv_counts[e]), float(n));
        expect_total += n;
      }
      if (num_recv[0] != expect_total) mismatch("ep num_recv", rank, call, 0, float(num_recv[0]), float(expect_total));
      int row = 0;
      for (int e = 0; e < epr; ++e)
        for (int i = 0; i < recv_counts[e]; ++i, ++row) {
          const int src = recv_src[row] / pairs, pair = recv_src[row] % pairs, t = pair / topk;
          if (src < 0 || src >= world || idval(call, src, t, pair % topk) != rank * epr + e) {
            mismatch("ep provenance", rank, call, row, float(recv_src[row]), float(e));
            continue;
          }
          for (int j = 0; j < hidden; ++j) {
            float got;
            if (fp8) {  // e4m3 payload + one scale per 128 values: small integers survive within the format's 3 mantissa bits
              const unsigned char b = static_cast<unsigned char>(recv_buf[size_t(row) * hidden + j]);
              const int ex = (b >> 3) & 15, m = b & 7;
              const float mag = ex == 0 ? ldexpf(float(m), -9) : ldexpf(float(8 + m), ex - 10);
              got = (b & 0x80 ? -mag : mag) * recv_scale[size_t(row) * (hidden / 128) + j / 128];
              if (fabsf(got - xval(call, src, t, j)) > 0.07f * 63.f) mismatch("ep fp8 payload", rank, call, long(row) * hidden + j, got, xval(call, src, t, j));
            } else {
              got = __bfloat162float(reinterpret_cast<const bf16*>(recv_buf)[size_t(row) * hidden + j]);
              if (got != xval(call, src, t, j)) mismatch("ep payload", rank, call, long(row) * hidden + j, got, xval(call, src, t, j));
            }
          }
        }
      if (!fp8) {
        // identity experts, then the way back
        memcpy(expert_out, recv_buf, size_t(num_recv[0]) * hidden * 2);
        bf16* op = out.data();
        sim::launch(rank, ctas, 128, [=] {
          tdist::ep_combine_kernel<tdist::BF16, bf16>(ctx, expert_out, idp, sip, static_cast<const float*>(nullptr), op, tokens, topk, hidden,
                                                      experts, flags + 32, gc + 2, ps);
        });
        for (int t = 0; t < tokens; ++t)
          for (int j = 0; j < hidden; ++j) {
            const float got = __bfloat162float(out[size_t(t) * hidden + j]), want = topk * xval(call, rank, t, j);
            if (got != want) mismatch("ep combine", rank, call, long(t) * hidden + j, got, want);
          }
      } else {
        // no combine in this mode: keep the ranks from starting the next dispatch while peers still check their buffers
        sim::launch(rank, 1, 32, [=] { tdist::barrier_all_kernel(ctx, flags + 48, ps); });
      }
    }
  });
  printf("protocol_sim ep: %d ranks x %d calls, %d tokens top-2 hidden %d over %d experts%s, %d CTAs: %s\n", world, calls, tokens, hidden,
         experts, fp8 ? ", fp8 payload" : "", ctas, g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

}  // namespace
