// mega_scenarios.h — the decode megakernel on the simulator: the REAL interpreter kernel (csrc/mega/megakernel.cu) runs the
// schedule the REAL native builder (csrc/mega/builder.cc) makes for a small Qwen3-style decoder, with 1-4 tensor-parallel
// ranks, several steps back to back without any host-side synchronisation between ranks, and a double-precision host model
// as the oracle. What the sanitizer sees here and nowhere else: the scoreboard (release increments, acquire polls) that
// orders every activation buffer between the CTAs of a GPU, the arrival counters of the fused attention task, the per-tile
// cross-rank flag exchange in front of every in-kernel all-reduce and the parity double buffering of its inputs across
// steps. The TMA weight stream (3-D tiled loads with the 128 B swizzle into per-warp stage buffers, mbarrier phases) and the
// ldmatrix / mma.sync fragments are emulated functionally (csrc/sim/shim/device/ptx.cuh). Included by protocol_sim.cc after
// the kernel sources; the builder itself is included at the very end of that file.
#pragma once

// the launch entry point the builder calls (the real one, behind the cut of megakernel.cu, uses the runtime API)
extern "C" int tdist_mega_launch(const void* params, int num_ctas, int smem_bytes, void*) {
  const tdist::mega::KernelParams p = *static_cast<const tdist::mega::KernelParams*>(params);
  if (smem_bytes > tdist::mega::kSmemLimit) {
    fprintf(stderr, "mega: %d bytes of dynamic shared memory requested\n", smem_bytes);
    return -1;
  }
  sim::launch(p.ctx.rank, num_ctas, tdist::mega::kThreads, [p] { tdist::mega::megakernel(p); });
  return 0;
}

namespace {

double bd(bf16 v) { return double(__bfloat162float(v)); }
double rb(double v) { return bd(__float2bfloat16(float(v))); }  // what a bf16 store keeps

struct MegaLayer {
  std::vector<bf16> in_norm, wqkv, qn, kn, wo, post_norm, gate_up, down;
  std::vector<bf16> kc, vc;  // paged cache [pages][page][Hkv_loc][128]
  std::vector<float> partial;  // split-KV partials of this layer's attention (+ the arrival counters of the one-task form)
};
struct MegaRank {
  std::vector<MegaLayer> L;
  std::vector<bf16> hidden, qkv, q_rope, attn_out, act, fc1_out, normed, tmp, logits;
  std::vector<long long> ids, next_ids;
  std::vector<int> positions;
};
struct MegaWant {  // per step
  std::vector<double> hidden, logits;
  std::vector<std::vector<std::vector<double>>> new_k, new_v;  // [rank][layer][B * Hkv_loc * 128]
};

std::vector<bf16> mega_rand(size_t n, unsigned seed, float scale, float offset = 0.f) {
  std::vector<bf16> v(n);
  for (auto& x : v) x = __float2bfloat16(offset + sim::unit_rand(seed) * scale);
  return v;
}

// graph: 0 = fused (norm in the linear prologue, SiLU * up in fc1, residual in the all-reduce), 1 = op by op as the
// reference records it (+ a cross-rank barrier op per layer), 2 = fused + the one-task attention
int run_mega(int world, int ctas, int layers, int B, int steps, int graph, int multicast, int kvl, int g, int inter_loc, int splits) {
  using namespace tdist::mega;
  constexpr int D = 128;
  const int H = 256, vocab = 320, page = 8, pos0 = 13;
  const int Hq = kvl * g, heads = Hq + 2 * kvl;
  const int max_pos = pos0 + 3 * B + steps + 1, max_pages = (max_pos + page - 1) / page, pages = B * max_pages;
  const float eps = 1e-6f, scale = 1.f / std::sqrt(float(D));
  if (B < 1 || B > 8 || inter_loc % 64 != 0 || world > TDIST_MAX_RANKS || (graph == 2 && g > 8)) {
    fprintf(stderr, "mega: need 1 <= batch <= 8, intermediate size per rank a multiple of 64, at most 8 query heads per KV head for the fused attention\n");
    return 2;
  }
  // ---- model and inputs (replicated tensors are shared by the ranks: they are only read)
  const auto embed = mega_rand(size_t(vocab) * H, 11, 1.f), final_norm = mega_rand(H, 12, 0.2f, 1.f);
  const auto lm_head = mega_rand(size_t(vocab) * H, 13, 2.f / std::sqrt(float(H)));
  std::vector<float> cos_sin(size_t(max_pos) * D);
  for (int p = 0; p < max_pos; ++p)
    for (int i = 0; i < 64; ++i) {
      const double ang = p * std::pow(10000.0, -double(i) / 64);
      cos_sin[size_t(p) * D + i] = float(std::cos(ang));
      cos_sin[size_t(p) * D + 64 + i] = float(std::sin(ang));
    }
  std::vector<int> block_table(size_t(B) * max_pages);
  for (int b = 0; b < B; ++b)
    for (int p = 0; p < max_pages; ++p) block_table[size_t(b) * max_pages + p] = (max_pages - 1 - p) * B + b;
  const size_t page_stride = size_t(page) * kvl * D, seq_stride = size_t(kvl) * D;
  auto cache_off = [&](int b, int pos, int kvh) {
    return size_t(block_table[size_t(b) * max_pages + pos / page]) * page_stride + size_t(pos % page) * seq_stride + size_t(kvh) * D;
  };
  std::vector<MegaRank> R(world);
  for (int r = 0; r < world; ++r) {
    MegaRank& m = R[r];
    m.L.resize(layers);
    for (int l = 0; l < layers; ++l) {
      MegaLayer& w = m.L[l];
      const unsigned s = 1000u * l, sr = s + 100u * (r + 1);
      w.in_norm = mega_rand(H, s + 1, 0.2f, 1.f);  // norm weights are replicated
      w.post_norm = mega_rand(H, s + 2, 0.2f, 1.f);
      w.qn = mega_rand(D, s + 3, 0.2f, 1.f);
      w.kn = mega_rand(D, s + 4, 0.2f, 1.f);
      w.wqkv = mega_rand(size_t(heads) * D * H, sr + 5, 1.5f / std::sqrt(float(H)));
      w.wo = mega_rand(size_t(H) * Hq * D, sr + 6, 1.5f / std::sqrt(float(world * Hq * D)));
      w.gate_up = mega_rand(size_t(2) * inter_loc * H, sr + 7, 1.5f / std::sqrt(float(H)));
      w.down = mega_rand(size_t(H) * inter_loc, sr + 8, 1.5f / std::sqrt(float(world * inter_loc)));
      w.kc = mega_rand(size_t(pages) * page_stride, sr + 9, 1.f);
      w.vc = mega_rand(size_t(pages) * page_stride, sr + 10, 1.f);
      w.partial.assign(size_t(B) * Hq * splits * (D + 1) + size_t(B) * kvl, 0.f);
    }
    m.hidden.resize(size_t(B) * H);
    m.qkv.resize(size_t(B) * heads * D);
    m.q_rope.resize(size_t(B) * Hq * D);
    m.attn_out.resize(size_t(B) * Hq * D);
    m.act.resize(size_t(B) * inter_loc);
    m.fc1_out.resize(size_t(B) * 2 * inter_loc);
    m.normed.resize(size_t(B) * H);
    m.tmp.resize(size_t(B) * H);
    m.logits.resize(size_t(B) * vocab);
    m.ids.resize(B);
    m.next_ids.assign(B, -1);
    m.positions.resize(B);
  }
  std::vector<std::vector<long long>> ids(steps, std::vector<long long>(B));
  unsigned seed = 777;
  for (auto& v : ids)
    for (auto& i : v) i = (lcg(seed) >> 8) % vocab;

  // ---- oracle: the whole model in double precision, rounded to bf16 wherever the kernels store bf16
  std::vector<MegaWant> want(steps);
  {
    std::vector<std::vector<std::vector<bf16>>> kc(world), vc(world);
    for (int r = 0; r < world; ++r)
      for (int l = 0; l < layers; ++l) {
        kc[r].push_back(R[r].L[l].kc);
        vc[r].push_back(R[r].L[l].vc);
      }
    auto rmsnorm = [&](const std::vector<double>& x, const std::vector<bf16>& w, std::vector<double>& y) {
      y.resize(x.size());
      for (int b = 0; b < B; ++b) {
        double ss = 0;
        for (int i = 0; i < H; ++i) ss += x[size_t(b) * H + i] * x[size_t(b) * H + i];
        const double inv = 1.0 / std::sqrt(ss / H + eps);
        for (int i = 0; i < H; ++i) y[size_t(b) * H + i] = rb(rb(x[size_t(b) * H + i] * inv) * bd(w[i]));
      }
    };
    auto linear = [&](const std::vector<double>& x, const std::vector<bf16>& w, int N, int K, std::vector<double>& y) {
      y.assign(size_t(B) * N, 0.0);
      for (int b = 0; b < B; ++b)
        for (int n = 0; n < N; ++n) {
          double acc = 0;
          for (int k = 0; k < K; ++k) acc += x[size_t(b) * K + k] * bd(w[size_t(n) * K + k]);
          y[size_t(b) * N + n] = rb(acc);
        }
    };
    for (int st = 0; st < steps; ++st) {
      MegaWant& W = want[st];
      W.new_k.assign(world, std::vector<std::vector<double>>(layers));
      W.new_v = W.new_k;
      std::vector<double> h(size_t(B) * H), xn, qkv, part, gu;
      for (int b = 0; b < B; ++b)
        for (int i = 0; i < H; ++i) h[size_t(b) * H + i] = bd(embed[size_t(ids[st][b]) * H + i]);
      for (int l = 0; l < layers; ++l) {
        std::vector<double> sum(size_t(B) * H, 0.0);
        for (int r = 0; r < world; ++r) {
          const MegaLayer& w = R[r].L[l];
          rmsnorm(h, w.in_norm, xn);
          linear(xn, w.wqkv, heads * D, H, qkv);
          std::vector<double> attn(size_t(B) * Hq * D);
          W.new_k[r][l].resize(size_t(B) * kvl * D);
          W.new_v[r][l].resize(size_t(B) * kvl * D);
          std::vector<double> q(size_t(B) * Hq * D);
          for (int b = 0; b < B; ++b) {
            const int pos = pos0 + 3 * b + st;
            for (int hd = 0; hd < heads; ++hd) {
              const double* src = qkv.data() + (size_t(b) * heads + hd) * D;
              double v[D], o[D];
              const bool is_q = hd < Hq, is_k = !is_q && hd < Hq + kvl;
              if (is_q || is_k) {
                double ss = 0;
                for (int i = 0; i < D; ++i) ss += src[i] * src[i];
                const double inv = 1.0 / std::sqrt(ss / D + eps);
                for (int i = 0; i < D; ++i) v[i] = rb(rb(src[i] * inv) * bd((is_q ? w.qn : w.kn)[i]));
                const float* cs = cos_sin.data() + size_t(pos) * D;
                for (int i = 0; i < 64; ++i) {
                  o[i] = rb(v[i] * cs[i] - v[i + 64] * cs[64 + i]);
                  o[i + 64] = rb(v[i + 64] * cs[i] + v[i] * cs[64 + i]);
                }
              } else {
                for (int i = 0; i < D; ++i) o[i] = src[i];
              }
              if (is_q) {
                for (int i = 0; i < D; ++i) q[(size_t(b) * Hq + hd) * D + i] = o[i];
              } else {
                const int kvh = is_k ? hd - Hq : hd - Hq - kvl;
                auto& cache = is_k ? kc[r][l] : vc[r][l];
                auto& rec = is_k ? W.new_k[r][l] : W.new_v[r][l];
                for (int i = 0; i < D; ++i) {
                  cache[cache_off(b, pos, kvh) + i] = __float2bfloat16(float(o[i]));
                  rec[(size_t(b) * kvl + kvh) * D + i] = o[i];
                }
              }
            }
            for (int hd = 0; hd < Hq; ++hd) {
              const int kvh = hd / g;
              std::vector<double> sc(pos + 1);
              double mx = -1e300, den = 0, o[D] = {};
              for (int p = 0; p <= pos; ++p) {
                double s = 0;
                const size_t off = cache_off(b, p, kvh);
                for (int i = 0; i < D; ++i) s += q[(size_t(b) * Hq + hd) * D + i] * bd(kc[r][l][off + i]);
                sc[p] = s * scale;
                mx = std::max(mx, sc[p]);
              }
              for (int p = 0; p <= pos; ++p) {
                const double e = std::exp(sc[p] - mx);
                den += e;
                const size_t off = cache_off(b, p, kvh);
                for (int i = 0; i < D; ++i) o[i] += e * bd(vc[r][l][off + i]);
              }
              for (int i = 0; i < D; ++i) attn[(size_t(b) * Hq + hd) * D + i] = rb(o[i] / den);
            }
          }
          linear(attn, w.wo, H, Hq * D, part);
          for (size_t i = 0; i < sum.size(); ++i) sum[i] += part[i];
        }
        for (size_t i = 0; i < h.size(); ++i) h[i] = rb(rb(sum[i]) + h[i]);
        std::fill(sum.begin(), sum.end(), 0.0);
        for (int r = 0; r < world; ++r) {
          const MegaLayer& w = R[r].L[l];
          rmsnorm(h, w.post_norm, xn);
          linear(xn, w.gate_up, 2 * inter_loc, H, gu);
          std::vector<double> act(size_t(B) * inter_loc);
          for (int b = 0; b < B; ++b)
            for (int i = 0; i < inter_loc; ++i) {
              const double gt = gu[size_t(b) * 2 * inter_loc + i], up = gu[size_t(b) * 2 * inter_loc + inter_loc + i];
              act[size_t(b) * inter_loc + i] = rb(gt / (1.0 + std::exp(-gt)) * up);
            }
          linear(act, w.down, H, inter_loc, part);
          for (size_t i = 0; i < sum.size(); ++i) sum[i] += part[i];
        }
        for (size_t i = 0; i < h.size(); ++i) h[i] = rb(rb(sum[i]) + h[i]);
      }
      W.hidden = h;
      rmsnorm(h, final_norm, xn);
      linear(xn, lm_head, vocab, H, W.logits);
    }
  }

  // ---- the ranks
  const size_t part_elems = size_t(2) * B * H;  // two parities
  const int cross_ops = layers * 3 + 1;
  const size_t flag_words = size_t(cross_ops) * kMaxArTiles * TDIST_MAX_RANKS;
  sim::Node node(world, 2 * part_elems * sizeof(bf16) + flag_words * sizeof(uint32_t), multicast != 0);
  std::atomic<int> tasks_total{0}, ops_total{0};
  sim::run_ranks(world, [&](int rank) {
    MegaRank& m = R[rank];
    bf16* o_partial = reinterpret_cast<bf16*>(node.base(rank));
    bf16* down_partial = o_partial + part_elems;
    uint32_t* ar_flags = reinterpret_cast<uint32_t*>(down_partial + part_elems);
    void* hb = nullptr;
    tdist_mega_create(rank, world, ctas, &hb);
    auto add = [&](int type, std::initializer_list<const void*> ins, std::initializer_list<void*> outs, int M, int N, int K,
                   std::initializer_list<long long> ints = {}, std::initializer_list<long long> longs = {},
                   std::initializer_list<float> floats = {}, int dep0 = -2, int dep1 = -1, int chain = 1) {
      tdist_mega_op d;
      memset(&d, 0, sizeof(d));
      int k = 0;
      for (const void* p : ins) d.in[k++] = p;
      k = 0;
      for (void* p : outs) d.out[k++] = p;
      d.M = M, d.N = N, d.K = K;
      k = 0;
      for (long long v : ints) d.i[k++] = int(v);
      k = 0;
      for (long long v : longs) d.l[k++] = v;
      k = 0;
      for (float v : floats) d.f[k++] = v;
      const int idx = tdist_mega_add_op(hb, type, &d, dep0, dep1, chain);
      if (idx < 0) {
        fprintf(stderr, "mega: add_op(%d) refused\n", type);
        g_errors.fetch_add(1);
      }
      return idx;
    };
    auto fc = [&](const bf16* x, const std::vector<bf16>& w, bf16* y, int N, int K, int ldy, const bf16* norm_w, bool pair, long long parity) {
      return add(OP_LINEAR, {x, w.data(), norm_w, nullptr}, {y}, B, N, K, {pair ? 1 : 0, ldy, parity}, {}, {eps});
    };
    const long long parity_bytes = (long long)B * H * sizeof(bf16);
    std::vector<std::vector<bf16>> qk_w(layers);  // [2][128] q / k norm weights of the one-task attention, per layer
    add(OP_EMBEDDING, {m.ids.data(), embed.data()}, {m.hidden.data()}, B, 0, H);
    for (int l = 0; l < layers; ++l) {
      MegaLayer& w = m.L[l];
      if (graph != 1) {
        fc(m.hidden.data(), w.wqkv, m.qkv.data(), heads * D, H, heads * D, w.in_norm.data(), false, 0);
      } else {
        add(OP_RMS_NORM, {m.hidden.data(), nullptr, w.in_norm.data(), nullptr}, {m.normed.data()}, B, 0, H, {}, {}, {eps});
        fc(m.normed.data(), w.wqkv, m.qkv.data(), heads * D, H, heads * D, nullptr, false, 0);
      }
      if (graph == 2) {
        qk_w[l] = w.qn;
        qk_w[l].insert(qk_w[l].end(), w.kn.begin(), w.kn.end());
        add(OP_ATTN_FUSED, {m.qkv.data(), w.kc.data(), w.vc.data(), block_table.data(), m.positions.data(), cos_sin.data()},
            {w.partial.data(), m.attn_out.data(), qk_w[l].data()}, B, 0, 0, {Hq, kvl, page, max_pages, splits},
            {(long long)page_stride, (long long)seq_stride, D}, {scale, eps});
      } else {
        add(OP_QK_NORM_ROPE_KV, {m.qkv.data(), w.kc.data(), w.vc.data(), w.qn.data(), w.kn.data(), cos_sin.data()},
            {m.q_rope.data(), m.positions.data(), block_table.data()}, B, 0, 0, {Hq, kvl, page, max_pages},
            {(long long)page_stride, (long long)seq_stride, D}, {eps});
        add(OP_ATTN_SPLIT, {m.q_rope.data(), w.kc.data(), w.vc.data(), block_table.data(), m.positions.data()}, {w.partial.data()}, B, 0, 0,
            {Hq, kvl, page, max_pages, splits}, {(long long)page_stride, (long long)seq_stride, D}, {scale});
        add(OP_ATTN_COMBINE, {w.partial.data()}, {m.attn_out.data()}, B, 0, 0, {Hq, kvl, page, max_pages, splits});
      }
      fc(m.attn_out.data(), w.wo, o_partial, H, Hq * D, H, nullptr, false, parity_bytes);
      if (graph != 1) {
        add(OP_ALLREDUCE, {o_partial, nullptr, nullptr, m.hidden.data()}, {nullptr}, 0, B * H, 0, {0, 0, parity_bytes});
        fc(m.hidden.data(), w.gate_up, m.act.data(), inter_loc, H, inter_loc, w.post_norm.data(), true, 0);
        fc(m.act.data(), w.down, down_partial, H, inter_loc, H, nullptr, false, parity_bytes);
        add(OP_ALLREDUCE, {down_partial, nullptr, nullptr, m.hidden.data()}, {nullptr}, 0, B * H, 0, {0, 0, parity_bytes});
      } else {
        add(OP_ALLREDUCE, {o_partial, nullptr, nullptr, nullptr}, {m.tmp.data()}, 0, B * H, 0, {0, 0, parity_bytes});
        add(OP_ADD, {m.hidden.data(), m.tmp.data()}, {m.hidden.data()}, 0, B * H, 0);
        add(OP_RMS_NORM, {m.hidden.data(), nullptr, w.post_norm.data(), nullptr}, {m.normed.data()}, B, 0, H, {}, {}, {eps});
        fc(m.normed.data(), w.gate_up, m.fc1_out.data(), 2 * inter_loc, H, 2 * inter_loc, nullptr, false, 0);
        add(OP_SILU_MUL, {m.fc1_out.data()}, {m.act.data()}, B, inter_loc, 0);
        fc(m.act.data(), w.down, down_partial, H, inter_loc, H, nullptr, false, parity_bytes);
        add(OP_ALLREDUCE, {down_partial, nullptr, nullptr, nullptr}, {m.tmp.data()}, 0, B * H, 0, {0, 0, parity_bytes});
        add(OP_ADD, {m.hidden.data(), m.tmp.data()}, {m.hidden.data()}, 0, B * H, 0);
        add(OP_BARRIER, {}, {}, 0, 0, 0);
      }
    }
    fc(m.hidden.data(), lm_head, m.logits.data(), vocab, H, vocab, final_norm.data(), false, 0);
    add(OP_ARGMAX, {m.logits.data()}, {m.next_ids.data()}, B, vocab, 0, {0, vocab});
    tdist_symm_ctx sc;
    memset(&sc, 0, sizeof(sc));
    const tdist::SymmCtx c = node.ctx(rank);
    for (int q = 0; q < TDIST_MAX_RANKS; ++q) sc.base[q] = c.base[q];
    sc.mc_base = c.mc_base;
    sc.bytes = c.bytes;
    sc.rank = rank;
    sc.world = world;
    if (tdist_mega_compile(hb, &sc, ar_flags) != 0) {
      fprintf(stderr, "mega: compile failed\n");
      g_errors.fetch_add(1);
      return;
    }
    int info[6];
    tdist_mega_info(hb, info);
    if (rank == 0) {
      ops_total = info[0];
      tasks_total = info[1];
    }
    auto close = [&](const char* what, int step, long i, double got, double wantv) {
      if (!(std::fabs(got - wantv) <= 0.03 * std::fabs(wantv) + 0.03)) mismatch(what, rank, step, i, float(got), float(wantv));
    };
    for (int st = 0; st < steps; ++st) {
      for (int b = 0; b < B; ++b) {
        m.ids[b] = ids[st][b];
        m.positions[b] = pos0 + 3 * b + st;
      }
      if (tdist_mega_run(hb, nullptr) != 0) {
        g_errors.fetch_add(1);
        return;
      }
      const MegaWant& W = want[st];
      for (size_t i = 0; i < m.hidden.size(); ++i) close("mega hidden", st, long(i), bd(m.hidden[i]), W.hidden[i]);
      for (size_t i = 0; i < m.logits.size(); ++i) close("mega logits", st, long(i), bd(m.logits[i]), W.logits[i]);
      for (int l = 0; l < layers; ++l)
        for (int b = 0; b < B; ++b)
          for (int kvh = 0; kvh < kvl; ++kvh)
            for (int i = 0; i < D; ++i) {
              const size_t off = cache_off(b, m.positions[b], kvh) + i, w = (size_t(b) * kvl + kvh) * D + i;
              close("mega K cache", st, long(off), bd(m.L[l].kc[off]), W.new_k[rank][l][w]);
              close("mega V cache", st, long(off), bd(m.L[l].vc[off]), W.new_v[rank][l][w]);
            }
      for (int b = 0; b < B; ++b) {  // greedy token: exact on the kernel's own logits, lowest index on ties
        int best = 0;
        for (int i = 1; i < vocab; ++i)
          if (bd(m.logits[size_t(b) * vocab + i]) > bd(m.logits[size_t(b) * vocab + best])) best = i;
        if (m.next_ids[b] != best) mismatch("mega argmax", rank, st, b, float(m.next_ids[b]), float(best));
      }
    }
    tdist_mega_destroy(hb);
  });
  printf("protocol_sim mega: %d rank(s) x %d CTAs, %d layer(s), batch %d, %d step(s), graph %s, %s all-reduce, %d KV / %d query heads and "
         "%d MLP columns per rank, %d KV splits: %d ops, %d tasks per step: %s\n",
         world, ctas, layers, B, steps, graph == 0 ? "fused" : graph == 1 ? "op by op" : "fused + one-task attention",
         multicast ? "in-switch" : "peer-load", kvl, Hq, inter_loc, splits, ops_total.load(), tasks_total.load(),
         g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

}  // namespace
