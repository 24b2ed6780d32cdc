// decode_scenarios.h — simulator scenario for the PRODUCTION decode-path kernels (csrc/kernels/decode.cu): one decode step
// of a dense transformer layer on one simulated GPU, every kernel checked against a host restatement in double precision
// computed from the kernel's own inputs. No communication: this is a functional check (indexing, paging, GQA passes, empty
// KV splits, ties in the sampler) of kernels that otherwise only run on a GPU; the sanitizer additionally sees every pair of
// CTAs that touch the same bytes. Included by protocol_sim.cc after the kernel sources.
#pragma once

namespace {

template <typename T>
struct Tol;
template <>
struct Tol<__nv_bfloat16> {
  static constexpr double rel = 1.0 / 96, abs = 2e-3;
  static constexpr const char* name = "bf16";
};
template <>
struct Tol<__half> {
  static constexpr double rel = 1.0 / 768, abs = 5e-4;
  static constexpr const char* name = "fp16";
};

template <typename T>
std::vector<T> random_16(size_t n, unsigned seed, float scale) {
  std::vector<T> v(n);
  for (auto& x : v) x = tdist::Cvt<T>::one(sim::unit_rand(seed) * scale);
  return v;
}
template <typename T>
double up(T v) { return double(tdist::Cvt<T>::f(v)); }
template <typename T>
void expect_close(const char* what, long i, T got, double want, double scale = 1.0) {
  const double g = up(got);
  if (!(std::fabs(g - want) <= Tol<T>::rel * std::fabs(want) + Tol<T>::abs * scale)) mismatch(what, 0, 0, i, float(g), float(want));
}

// y[m, n] = sum_k x[m, k] w[n, k] + bias[n], launched as tdist_gemv launches it
template <typename T>
void sim_gemv(const char* what, const std::vector<T>& x, const std::vector<T>& w, const T* bias, std::vector<T>& y, int M, int N, int K) {
  const int grid = std::min((N + 15) / 16, 148 * 8);
  const T *xp = x.data(), *wp = w.data();
  T* yp = y.data();
  auto go = [&](auto maxm) {
    constexpr int MAXM = decltype(maxm)::value;
    sim::launch(0, grid, 256, [=] { tdist::gemv_kernel<T, MAXM>(xp, wp, bias, yp, M, N, K, K, K, N); });
  };
  if (M <= 1) go(std::integral_constant<int, 1>{});
  else if (M <= 2) go(std::integral_constant<int, 2>{});
  else if (M <= 4) go(std::integral_constant<int, 4>{});
  else if (M <= 8) go(std::integral_constant<int, 8>{});
  else go(std::integral_constant<int, 16>{});
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double acc = bias ? up(bias[n]) : 0.0, mag = 0.0;
      for (int k = 0; k < K; ++k) {
        const double p = up(x[size_t(m) * K + k]) * up(w[size_t(n) * K + k]);
        acc += p;
        mag += std::fabs(p);
      }
      expect_close(what, long(m) * N + n, y[size_t(m) * N + n], acc, 1.0 + mag * 1e-3);
    }
}

template <typename T>
int run_decode_t(int B, int Hq, int Hkv, int kv_max, int splits, int page_size, int hidden, int inter, int vocab) {
  constexpr int D = 128;
  const int heads = Hq + 2 * Hkv, g_total = Hq / Hkv;
  const float eps = 1e-5f, sm_scale = 1.f / std::sqrt(float(D));
  unsigned seed = 4242;

  // ---- embedding
  std::vector<long long> ids(B);
  for (auto& i : ids) i = (lcg(seed) >> 8) % vocab;
  const auto table = random_16<T>(size_t(vocab) * hidden, 1, 1.f);
  std::vector<T> x(size_t(B) * hidden);
  sim::launch(0, B, 256, [&] { tdist::embedding_kernel<T>(ids.data(), table.data(), x.data(), hidden); });
  for (int b = 0; b < B; ++b)
    for (int i = 0; i < hidden; ++i)
      if (up(x[size_t(b) * hidden + i]) != up(table[size_t(ids[b]) * hidden + i]))
        mismatch("embedding", 0, 0, long(b) * hidden + i, float(up(x[size_t(b) * hidden + i])), float(up(table[size_t(ids[b]) * hidden + i])));

  // ---- RMSNorm with the fused residual add (residual updated in place), then without a residual
  auto resid = random_16<T>(size_t(B) * hidden, 2, 1.f);
  const auto resid_in = resid;
  const auto norm_w = random_16<T>(hidden, 3, 1.f);
  std::vector<T> xn(size_t(B) * hidden), xn_plain(size_t(B) * hidden);
  sim::launch(0, B, 256, [&] { tdist::rmsnorm_kernel<T>(x.data(), resid.data(), norm_w.data(), xn.data(), hidden, eps, hidden, hidden); });
  sim::launch(0, B, 256, [&] { tdist::rmsnorm_kernel<T>(x.data(), static_cast<T*>(nullptr), norm_w.data(), xn_plain.data(), hidden, eps, hidden, hidden); });
  for (int b = 0; b < B; ++b) {
    double ss = 0, ss_plain = 0;
    for (int i = 0; i < hidden; ++i) {
      const size_t k = size_t(b) * hidden + i;
      expect_close("rmsnorm residual", long(k), resid[k], up(x[k]) + up(resid_in[k]));
      ss += up(resid[k]) * up(resid[k]);  // the norm sees the ROUNDED sum
      ss_plain += up(x[k]) * up(x[k]);
    }
    const double inv = 1.0 / std::sqrt(ss / hidden + eps), inv_plain = 1.0 / std::sqrt(ss_plain / hidden + eps);
    for (int i = 0; i < hidden; ++i) {
      const size_t k = size_t(b) * hidden + i;
      expect_close("rmsnorm", long(k), xn[k], up(resid[k]) * inv * up(norm_w[i]));
      expect_close("rmsnorm (no residual)", long(k), xn_plain[k], up(x[k]) * inv_plain * up(norm_w[i]));
    }
  }

  // ---- QKV projection (with bias)
  const auto w_qkv = random_16<T>(size_t(heads) * D * hidden, 4, 0.08f);
  const auto b_qkv = random_16<T>(size_t(heads) * D, 5, 0.5f);
  std::vector<T> qkv(size_t(B) * heads * D);
  sim_gemv<T>("gemv qkv", xn, w_qkv, b_qkv.data(), qkv, B, heads * D, hidden);

  // ---- q/k RMSNorm + rotary embedding + KV-cache append; the cache is dense [B, S, Hkv, 128] or paged
  // [pages, page_size, Hkv, 128] with the pages of a sequence scattered through the pool
  std::vector<int> kv_lens(B), positions(B), batch_idx(B);
  for (int b = 0; b < B; ++b) {
    kv_lens[b] = b == 0 ? kv_max : b == 1 ? std::min(kv_max, std::max(1, splits - 2)) : 1 + int((lcg(seed) >> 8) % kv_max);
    positions[b] = kv_lens[b] - 1;  // the new token is the last position the attention reads
    batch_idx[b] = b;
  }
  const int S = kv_max, max_pages = page_size > 0 ? (S + page_size - 1) / page_size : 0;
  const long long head_stride = D, seq_stride = (long long)Hkv * D;
  const long long batch_stride = page_size > 0 ? (long long)page_size * seq_stride : (long long)S * seq_stride;
  std::vector<int> block_table(size_t(B) * std::max(max_pages, 1));
  for (int b = 0; b < B; ++b)
    for (int p = 0; p < max_pages; ++p) block_table[size_t(b) * max_pages + p] = (max_pages - 1 - p) * B + b;  // reversed, interleaved
  const int* bt = page_size > 0 ? block_table.data() : nullptr;
  const size_t cache_elems = page_size > 0 ? size_t(B) * max_pages * batch_stride : size_t(B) * batch_stride;
  auto k_cache = random_16<T>(cache_elems, 6, 1.f), v_cache = random_16<T>(cache_elems, 7, 1.f);
  const auto k_before = k_cache, v_before = v_cache;
  auto cache_off = [&](int b, int pos, int kvh) -> size_t {
    if (page_size > 0) return size_t(block_table[size_t(b) * max_pages + pos / page_size]) * batch_stride + size_t(pos % page_size) * seq_stride + size_t(kvh) * head_stride;
    return size_t(b) * batch_stride + size_t(pos) * seq_stride + size_t(kvh) * head_stride;
  };
  std::vector<float> cos_sin(size_t(S) * D);
  for (int p = 0; p < S; ++p)
    for (int i = 0; i < 64; ++i) {
      const double ang = p * std::pow(10000.0, -double(i) / 64);
      cos_sin[size_t(p) * D + i] = float(std::cos(ang));
      cos_sin[size_t(p) * D + 64 + i] = float(std::sin(ang));
    }
  const auto qn_w = random_16<T>(D, 8, 1.f), kn_w = random_16<T>(D, 9, 1.f);
  std::vector<T> q(size_t(B) * Hq * D);
  const int rope_grid = (B * heads * 32 + 255) / 256;
  sim::launch(0, rope_grid, 256, [&] {
    tdist::qk_norm_rope_kv_kernel<T>(qkv.data(), qn_w.data(), kn_w.data(), cos_sin.data(), positions.data(), batch_idx.data(), q.data(), k_cache.data(),
                                     v_cache.data(), B, Hq, Hkv, eps, batch_stride, seq_stride, head_stride, bt, page_size, max_pages);
  });
  std::vector<char> written(cache_elems, 0);
  for (int b = 0; b < B; ++b)
    for (int h = 0; h < heads; ++h) {
      const T* src = qkv.data() + (size_t(b) * heads + h) * D;
      double v[D], o[D];
      for (int i = 0; i < D; ++i) v[i] = up(src[i]);
      const bool is_q = h < Hq, is_k = !is_q && h < Hq + Hkv;
      if (is_q || is_k) {
        double ss = 0;
        for (int i = 0; i < D; ++i) ss += v[i] * v[i];
        const double inv = 1.0 / std::sqrt(ss / D + eps);
        for (int i = 0; i < D; ++i) v[i] *= inv * up((is_q ? qn_w : kn_w)[i]);
        const float* cs = cos_sin.data() + size_t(positions[b]) * D;
        for (int i = 0; i < 64; ++i) {
          o[i] = v[i] * cs[i] - v[i + 64] * cs[64 + i];
          o[i + 64] = v[i + 64] * cs[i] + v[i] * cs[64 + i];
        }
      } else {
        for (int i = 0; i < D; ++i) o[i] = v[i];
      }
      if (is_q) {
        for (int i = 0; i < D; ++i) expect_close("rope q", long(b * Hq + h) * D + i, q[(size_t(b) * Hq + h) * D + i], o[i]);
      } else {
        const int kvh = is_k ? h - Hq : h - Hq - Hkv;
        const size_t off = cache_off(b, positions[b], kvh);
        for (int i = 0; i < D; ++i) {
          expect_close(is_k ? "rope k -> cache" : "v -> cache", long(off) + i, (is_k ? k_cache : v_cache)[off + i], o[i]);
          if (is_k) written[off + i] = 1;
        }
      }
    }
  for (size_t i = 0; i < cache_elems; ++i)  // nothing but the new position of each sequence may change
    if (!written[i] && (up(k_cache[i]) != up(k_before[i]) || up(v_cache[i]) != up(v_before[i])))
      mismatch("KV cache: a position other than the new one changed", 0, 0, long(i), float(up(k_cache[i])), float(up(k_before[i])));

  // ---- split-KV GQA decode attention: the head group of a KV head is covered in passes of 8 / 4 / 2 / 1 heads (as
  // launch_flash_decode does), then the log-sum-exp combine; a sequence shorter than `splits` leaves empty splits behind
  std::vector<float> partial(size_t(B) * Hq * splits * (D + 1), -77.f), out_lse(size_t(B) * Hq * (D + 1), -77.f);
  std::vector<T> attn(size_t(B) * Hq * D);
  auto fd_pass = [&](auto gc, int g_off) {
    constexpr int G = decltype(gc)::value;
    sim::launch3(splits, Hkv, B, 128, [&, g_off] {
      tdist::flash_decode_split_kernel<T, G>(q.data(), k_cache.data(), v_cache.data(), kv_lens.data(), partial.data(), Hq, Hkv, splits, sm_scale,
                                             batch_stride, seq_stride, head_stride, bt, page_size, max_pages, g_total, g_off);
    });
  };
  for (int done = 0; done < g_total;) {
    const int rem = g_total - done;
    if (rem >= 8) fd_pass(std::integral_constant<int, 8>{}, done), done += 8;
    else if (rem >= 4) fd_pass(std::integral_constant<int, 4>{}, done), done += 4;
    else if (rem >= 2) fd_pass(std::integral_constant<int, 2>{}, done), done += 2;
    else fd_pass(std::integral_constant<int, 1>{}, done), done += 1;
  }
  sim::launch3(Hq, B, 1, 128, [&] {
    tdist::flash_decode_combine_kernel<T>(partial.data(), attn.data(), out_lse.data(), Hq, splits, (long long)Hq * splits * (D + 1),
                                          (long long)splits * (D + 1), D + 1);
  });
  for (int b = 0; b < B; ++b)
    for (int h = 0; h < Hq; ++h) {
      const int kvh = h / g_total;
      std::vector<double> sc(kv_lens[b]);
      double mx = -1e300;
      for (int p = 0; p < kv_lens[b]; ++p) {
        double s = 0;
        const size_t off = cache_off(b, p, kvh);
        for (int i = 0; i < D; ++i) s += up(q[(size_t(b) * Hq + h) * D + i]) * up(k_cache[off + i]);
        sc[p] = s * sm_scale;
        mx = std::max(mx, sc[p]);
      }
      double den = 0, o[D] = {};
      for (int p = 0; p < kv_lens[b]; ++p) {
        const double w = std::exp(sc[p] - mx);
        den += w;
        const size_t off = cache_off(b, p, kvh);
        for (int i = 0; i < D; ++i) o[i] += w * up(v_cache[off + i]);
      }
      const float* lse_row = out_lse.data() + (size_t(b) * Hq + h) * (D + 1);
      for (int i = 0; i < D; ++i) {
        expect_close("flash decode", long(b * Hq + h) * D + i, attn[(size_t(b) * Hq + h) * D + i], o[i] / den);
        if (std::fabs(lse_row[i] - o[i] / den) > 1e-4 + 1e-4 * std::fabs(o[i] / den)) mismatch("flash decode fp32 out", 0, 0, long(b * Hq + h) * D + i, lse_row[i], float(o[i] / den));
      }
      const double lse = mx + std::log(den);
      if (std::fabs(lse_row[D] - lse) > 1e-3) mismatch("flash decode lse", 0, 0, long(b) * Hq + h, lse_row[D], float(lse));
    }

  // ---- output projection, residual add, gated MLP
  const auto w_o = random_16<T>(size_t(hidden) * Hq * D, 10, 0.08f);
  std::vector<T> proj(size_t(B) * hidden), h1(size_t(B) * hidden);
  sim_gemv<T>("gemv o_proj", attn, w_o, static_cast<const T*>(nullptr), proj, B, hidden, Hq * D);
  const long long add_vecs = (long long)B * hidden / 8;
  sim::launch(0, int(std::max<long long>(1, std::min<long long>((add_vecs + 255) / 256, 148 * 8))), 256,
              [&] { tdist::add_kernel<T>(proj.data(), resid.data(), h1.data(), add_vecs); });
  for (size_t i = 0; i < h1.size(); ++i) expect_close("add", long(i), h1[i], up(proj[i]) + up(resid[i]));
  const auto w_gu = random_16<T>(size_t(2) * inter * hidden, 11, 0.1f), w_down = random_16<T>(size_t(hidden) * inter, 12, 0.1f);
  std::vector<T> gate_up(size_t(B) * 2 * inter), act(size_t(B) * inter), down(size_t(B) * hidden);
  sim_gemv<T>("gemv gate_up", h1, w_gu, static_cast<const T*>(nullptr), gate_up, B, 2 * inter, hidden);
  const long long act_vecs = (long long)B * (inter / 8);
  sim::launch(0, int(std::max<long long>(1, std::min<long long>((act_vecs + 255) / 256, 148 * 8))), 256,
              [&] { tdist::silu_mul_kernel<T>(gate_up.data(), act.data(), inter, act_vecs); });
  for (int b = 0; b < B; ++b)
    for (int i = 0; i < inter; ++i) {
      const double g = up(gate_up[size_t(b) * 2 * inter + i]), u = up(gate_up[size_t(b) * 2 * inter + inter + i]);
      expect_close("silu_mul", long(b) * inter + i, act[size_t(b) * inter + i], g / (1.0 + std::exp(-g)) * u);
    }
  sim_gemv<T>("gemv down", act, w_down, static_cast<const T*>(nullptr), down, B, hidden, inter);

  // ---- LM head and greedy sampling; one extra row has its maximum three times: the lowest index wins
  const auto w_lm = random_16<T>(size_t(vocab) * hidden, 13, 0.2f);
  std::vector<T> logits(size_t(B + 1) * vocab);
  {
    std::vector<T> head(size_t(B) * vocab);
    sim_gemv<T>("gemv lm_head", down, w_lm, static_cast<const T*>(nullptr), head, B, vocab, hidden);
    std::copy(head.begin(), head.end(), logits.begin());
  }
  for (int i = 0; i < vocab; ++i) logits[size_t(B) * vocab + i] = tdist::Cvt<T>::one(float(i % 7) - 3.f);
  const int tie[3] = {vocab - 1, vocab / 3, vocab / 2};
  for (int t : tie) logits[size_t(B) * vocab + t] = tdist::Cvt<T>::one(9.f);
  std::vector<long long> next(B + 1, -1);
  sim::launch(0, B + 1, 1024, [&] { tdist::argmax_kernel<T>(logits.data(), next.data(), vocab, vocab); });
  for (int b = 0; b <= B; ++b) {
    int best = 0;
    for (int i = 1; i < vocab; ++i)
      if (up(logits[size_t(b) * vocab + i]) > up(logits[size_t(b) * vocab + best])) best = i;
    if (next[b] != best) mismatch("argmax", 0, 0, b, float(next[b]), float(best));
  }

  printf("protocol_sim decode (%s): batch %d, %d query / %d KV heads, up to %d cached tokens in %d splits, %s cache, hidden %d, "
         "intermediate %d, vocabulary %d: %s\n",
         Tol<T>::name, B, Hq, Hkv, kv_max, splits, page_size > 0 ? "paged" : "dense", hidden, inter, vocab, g_errors ? "DATA MISMATCH" : "data ok");
  return g_errors ? 1 : 0;
}

int run_decode(int B, int Hq, int Hkv, int kv_max, int splits, int page_size, int hidden, int inter, int vocab, int fp16) {
  if (Hq % Hkv != 0 || hidden % 8 != 0 || inter % 8 != 0 || B < 1 || B > 16) {
    fprintf(stderr, "decode: need Hq %% Hkv == 0, hidden and intermediate multiples of 8, 1 <= batch <= 16\n");
    return 2;
  }
  return fp16 ? run_decode_t<__half>(B, Hq, Hkv, kv_max, splits, page_size, hidden, inter, vocab)
              : run_decode_t<__nv_bfloat16>(B, Hq, Hkv, kv_max, splits, page_size, hidden, inter, vocab);
}

}  // namespace
