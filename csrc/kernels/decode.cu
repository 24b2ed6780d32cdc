// decode.cu — bandwidth-bound kernels of the decode path: weight-streaming
// GEMV, RMSNorm (+residual), per-head q/k RMSNorm + RoPE + KV-cache append,
// SiLU*up, GQA split-KV flash-decode and the log-sum-exp combine (intra- and
// inter-rank), embedding gather, greedy sampling.
//
// Reference counterparts: kernels/nvidia/flash_decode.py A1/K51-K53 (:129-532),
// mega_triton_kernel/kernels/{norm.py:32-227, activation.py:64, linear.py:32-99,
// flash_attn.py:32-163}; in the reference's layers the norm/rope/sampling steps
// come from flashinfer (layers/nvidia/tp_attn.py:29-43) and the linears from
// cuBLAS — here they are native so that no library kernel sits on the decode path.
#include <math_constants.h>

#include "common.cuh"

namespace tdist {

template <typename T>
struct Cvt;
template <>
struct Cvt<__nv_bfloat16> {
  __device__ static float2 up(uint32_t u) { return ptx::unpack_bf16x2(u); }
  __device__ static uint32_t down(float a, float b) { return ptx::pack_bf16x2(a, b); }
  __device__ static __nv_bfloat16 one(float a) { return __float2bfloat16(a); }
  __device__ static float f(__nv_bfloat16 a) { return __bfloat162float(a); }
};
template <>
struct Cvt<__half> {
  __device__ static float2 up(uint32_t u) { return ptx::unpack_f16x2(u); }
  __device__ static uint32_t down(float a, float b) { return ptx::pack_f16x2(a, b); }
  __device__ static __half one(float a) { return __float2half(a); }
  __device__ static float f(__half a) { return __half2float(a); }
};

template <typename T>
__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = Cvt<T>::up(u[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <typename T>
__device__ __forceinline__ int4 pack8(const float (&f)[8]) {
  int4 r;
  r.x = Cvt<T>::down(f[0], f[1]);
  r.y = Cvt<T>::down(f[2], f[3]);
  r.z = Cvt<T>::down(f[4], f[5]);
  r.w = Cvt<T>::down(f[6], f[7]);
  return r;
}

__device__ __forceinline__ float block_reduce_sum(float v, float* sh) {
  v = ptx::warp_reduce_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  if (l == 0) sh[w] = v;
  __syncthreads();
  float r = l < nw ? sh[l] : 0.f;
  r = ptx::warp_reduce_sum(r);
  __syncthreads();
  return r;
}

// ------------------------------------------------------------------ GEMV / skinny linear (M <= 16)
// y[m, n] = sum_k x[m, k] * w[n, k] (+ bias[n]). Each warp owns TWO output columns at a time and keeps 2 x 4
// sixteen-byte weight loads per lane in flight (4 KB per warp): with 8 warps per CTA and several CTAs per SM that is
// enough outstanding bytes to cover HBM latency (the one-column version reached 43% of peak, see profiles/).
// x is re-read through L1/L2 (it is tiny and hot). fp32 accumulate.
template <typename T, int MAXM>
__global__ void __launch_bounds__(256) gemv_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                   const T* __restrict__ bias, T* __restrict__ y, int M, int N, int K,
                                                   long long ldx, long long ldw, long long ldy) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  constexpr int UN = 4;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warps_per_cta = blockDim.x >> 5;
  const int kv = K >> 3;  // 16-byte vectors per row
  for (int n0 = (blockIdx.x * warps_per_cta + warp) * 2; n0 < N; n0 += gridDim.x * warps_per_cta * 2) {
    const bool two = n0 + 1 < N;
    float acc0[MAXM], acc1[MAXM];
#pragma unroll
    for (int m = 0; m < MAXM; ++m) acc0[m] = acc1[m] = 0.f;
    const T* w0 = w + size_t(n0) * ldw;
    const T* w1 = w + size_t(two ? n0 + 1 : n0) * ldw;
    for (int k0 = lane; k0 < kv; k0 += 32 * UN) {
      int4 a[UN], b[UN];
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        const int kk = k0 + u * 32;
        const bool ok = kk < kv;
        a[u] = ok ? ptx::ld_nc_v4(w0 + size_t(kk) * 8) : make_int4(0, 0, 0, 0);
        b[u] = ok ? ptx::ld_nc_v4(w1 + size_t(kk) * 8) : make_int4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        const int kk = k0 + u * 32;
        if (kk >= kv) break;
        float af[8], bf[8];
        unpack8<T>(a[u], af);
        unpack8<T>(b[u], bf);
#pragma unroll
        for (int m = 0; m < MAXM; ++m) {
          if (m < M) {
            float xf[8];
            unpack8<T>(*reinterpret_cast<const int4*>(x + size_t(m) * ldx + size_t(kk) * 8), xf);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              acc0[m] = fmaf(af[i], xf[i], acc0[m]);
              acc1[m] = fmaf(bf[i], xf[i], acc1[m]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int m = 0; m < MAXM; ++m) {
      if (m < M) {
        const float s0 = ptx::warp_reduce_sum(acc0[m]);
        const float s1 = ptx::warp_reduce_sum(acc1[m]);
        if (lane == 0) {
          y[size_t(m) * ldy + n0] = Cvt<T>::one(s0 + (bias ? Cvt<T>::f(bias[n0]) : 0.f));
          if (two) y[size_t(m) * ldy + n0 + 1] = Cvt<T>::one(s1 + (bias ? Cvt<T>::f(bias[n0 + 1]) : 0.f));
        }
      }
    }
  }
}

// ------------------------------------------------------------------ RMSNorm (+ fused residual add)
// y = rmsnorm(x (+ residual)) * weight ; if residual != nullptr it is updated in place to x + residual.
template <typename T>
__global__ void __launch_bounds__(256) rmsnorm_kernel(const T* __restrict__ x, T* residual, const T* __restrict__ weight,
                                                      T* __restrict__ y, int hidden, float eps, long long ldx,
                                                      long long ldy) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  extern __shared__ float sh_row[];  // hidden floats + 32 scratch
  float* scratch = sh_row + hidden;
  const size_t row = blockIdx.x;
  const int nv = hidden >> 3;
  float ss = 0.f;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    float f[8];
    unpack8<T>(*reinterpret_cast<const int4*>(x + row * ldx + size_t(i) * 8), f);
    if (residual) {
      float r[8];
      int4* rp = reinterpret_cast<int4*>(residual + row * hidden + size_t(i) * 8);
      unpack8<T>(*rp, r);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += r[j];
      *rp = pack8<T>(f);
      // the normalised value is computed from the rounded sum, like the unfused sequence would
      unpack8<T>(*rp, f);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sh_row[i * 8 + j] = f[j];
      ss = fmaf(f[j], f[j], ss);
    }
  }
  const float tot = block_reduce_sum(ss, scratch);
  const float inv = rsqrtf(tot / hidden + eps);
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    float wv[8], o[8];
    unpack8<T>(*reinterpret_cast<const int4*>(weight + size_t(i) * 8), wv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = sh_row[i * 8 + j] * inv * wv[j];
    *reinterpret_cast<int4*>(y + row * ldy + size_t(i) * 8) = pack8<T>(o);
  }
}

// ------------------------------------------------------------------ SiLU(gate) * up
// gate_up: [T, 2*I] (gate | up) -> out [T, I]
template <typename T>
__global__ void __launch_bounds__(256) silu_mul_kernel(const T* __restrict__ gate_up, T* __restrict__ out, int inter,
                                                       long long total_vec) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  const int iv = inter >> 3;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total_vec;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / iv;
    const int c = int(idx % iv);
    float g[8], u[8], o[8];
    unpack8<T>(*reinterpret_cast<const int4*>(gate_up + t * 2 * inter + size_t(c) * 8), g);
    unpack8<T>(*reinterpret_cast<const int4*>(gate_up + t * 2 * inter + inter + size_t(c) * 8), u);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = g[j] / (1.f + __expf(-g[j])) * u[j];
    *reinterpret_cast<int4*>(out + t * inter + size_t(c) * 8) = pack8<T>(o);
  }
}

// ------------------------------------------------------------------ residual add
template <typename T>
__global__ void __launch_bounds__(256) add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out,
                                                  long long n_vec) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_vec; i += (long long)gridDim.x * blockDim.x) {
    float x[8], y[8];
    unpack8<T>(reinterpret_cast<const int4*>(a)[i], x);
    unpack8<T>(reinterpret_cast<const int4*>(b)[i], y);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] += y[j];
    reinterpret_cast<int4*>(out)[i] = pack8<T>(x);
  }
}

// ------------------------------------------------------------------ q/k RMSNorm + RoPE + KV-cache append (head_dim 128)
// qkv: [T, (Hq + 2*Hkv) * 128]. One warp per (token, head). q heads are written to q_out [T, Hq, 128];
// k (normed + rotated) and v go to the cache at position positions[t] of sequence batch_idx[t].
// NeoX-style rotation: dim i pairs with i + 64. cos_sin: [max_pos, 128] = cos(64) | sin(64), fp32.
template <typename T>
__global__ void __launch_bounds__(256) qk_norm_rope_kv_kernel(
    const T* __restrict__ qkv, const T* __restrict__ q_norm_w, const T* __restrict__ k_norm_w,
    const float* __restrict__ cos_sin, const int* __restrict__ positions, const int* __restrict__ batch_idx,
    T* __restrict__ q_out, T* k_cache, T* v_cache, int T_tokens, int Hq, int Hkv, float eps, long long kv_batch_stride,
    long long kv_seq_stride, long long kv_head_stride, const int* __restrict__ block_table, int page_size,
    int max_pages) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int heads = Hq + 2 * Hkv;
  if (warp >= T_tokens * heads) return;
  const int t = warp / heads, h = warp % heads;
  const T* src = qkv + (size_t(t) * heads + h) * 128 + lane * 4;
  const uint2 raw = *reinterpret_cast<const uint2*>(src);
  float f[4];
  {
    const float2 a = Cvt<T>::up(raw.x), b = Cvt<T>::up(raw.y);
    f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y;
  }
  const bool is_q = h < Hq, is_k = !is_q && h < Hq + Hkv;
  const int pos = positions[t];
  if (is_q || is_k) {
    const T* nw = is_q ? q_norm_w : k_norm_w;
    if (nw != nullptr) {
      float ss = f[0] * f[0] + f[1] * f[1] + f[2] * f[2] + f[3] * f[3];
      ss = ptx::warp_reduce_sum(ss);
      const float inv = rsqrtf(ss / 128.f + eps);
      const uint2 wr = *reinterpret_cast<const uint2*>(nw + lane * 4);
      const float2 wa = Cvt<T>::up(wr.x), wb = Cvt<T>::up(wr.y);
      f[0] *= inv * wa.x, f[1] *= inv * wa.y, f[2] *= inv * wb.x, f[3] *= inv * wb.y;
    }
    // lanes 0-15 hold dims [0,64), lanes 16-31 hold dims [64,128): the rotation partner is lane ^ 16
    const float* cs = cos_sin + size_t(pos) * 128 + (lane & 15) * 4;
    const float4 c = *reinterpret_cast<const float4*>(cs), s = *reinterpret_cast<const float4*>(cs + 64);
    float o[4];
    const float cc[4] = {c.x, c.y, c.z, c.w}, sn[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float other = __shfl_xor_sync(0xffffffffu, f[i], 16);
      o[i] = lane < 16 ? f[i] * cc[i] - other * sn[i] : f[i] * cc[i] + other * sn[i];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) f[i] = o[i];
  }
  uint2 packed;
  packed.x = Cvt<T>::down(f[0], f[1]);
  packed.y = Cvt<T>::down(f[2], f[3]);
  if (is_q) {
    *reinterpret_cast<uint2*>(q_out + (size_t(t) * Hq + h) * 128 + lane * 4) = packed;
  } else {
    const int kvh = is_k ? h - Hq : h - Hq - Hkv;
    const int b = batch_idx ? batch_idx[t] : 0;
    size_t off;
    if (block_table != nullptr) {
      const int page = block_table[size_t(b) * max_pages + pos / page_size];
      off = size_t(page) * kv_batch_stride + size_t(pos % page_size) * kv_seq_stride + size_t(kvh) * kv_head_stride;
    } else {
      off = size_t(b) * kv_batch_stride + size_t(pos) * kv_seq_stride + size_t(kvh) * kv_head_stride;
    }
    T* dst = (is_k ? k_cache : v_cache) + off + lane * 4;
    *reinterpret_cast<uint2*>(dst) = packed;
  }
}

// ------------------------------------------------------------------ GQA split-KV flash decode (head_dim 128)
// grid (splits, Hkv, B); 128 threads. Half-warps (16 lanes x 8 dims) each own one KV position per
// step; every lane keeps the running softmax state of G query heads and 8 output dims per head.
// partial: [B, Hq, splits, 129] fp32 = unnormalised-then-normalised out (128) | lse (1).
template <typename T, int G>
__global__ void __launch_bounds__(128) flash_decode_split_kernel(
    const T* __restrict__ q, const T* __restrict__ k_cache, const T* __restrict__ v_cache,
    const int* __restrict__ kv_lens, float* __restrict__ partial, int Hq, int Hkv, int splits, float sm_scale,
    long long kv_batch_stride, long long kv_seq_stride, long long kv_head_stride, const int* __restrict__ block_table,
    int page_size, int max_pages, int g_total, int g_offset) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  constexpr int D = 128;
  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t half_mask = lane < 16 ? 0x0000ffffu : 0xffff0000u;
  const int sub = lane & 15;         // which 8-dim slice
  const int grp = warp * 2 + (lane >> 4);  // position group 0..7
  const int kv_len = kv_lens[b];
  const int per_split = (kv_len + splits - 1) / splits;
  const int s_beg = split * per_split, s_end = min(s_beg + per_split, kv_len);
  const int h0 = kvh * g_total + g_offset;  // first query head handled by this CTA pass

  float qf[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    unpack8<T>(*reinterpret_cast<const int4*>(q + (size_t(b) * Hq + h0 + g) * D + sub * 8), qf[g]);
#pragma unroll
    for (int i = 0; i < 8; ++i) qf[g][i] *= sm_scale * 1.4426950408889634f;  // scores in log2 units
  }
  float m[G], l[G], o[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -CUDART_INF_F;
    l[g] = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) o[g][i] = 0.f;
  }
  auto kv_off = [&](int pos) -> size_t {
    if (block_table != nullptr) {
      const int page = block_table[size_t(b) * max_pages + pos / page_size];
      return size_t(page) * kv_batch_stride + size_t(pos % page_size) * kv_seq_stride + size_t(kvh) * kv_head_stride;
    }
    return size_t(b) * kv_batch_stride + size_t(pos) * kv_seq_stride + size_t(kvh) * kv_head_stride;
  };
  constexpr int UN = 4;  // positions in flight per half-warp
  for (int p0 = s_beg + grp * UN; p0 < s_end; p0 += 8 * UN) {
    int4 kr[UN], vr[UN];
#pragma unroll
    for (int u = 0; u < UN; ++u) {
      const int pos = p0 + u;
      if (pos < s_end) {
        const size_t off = kv_off(pos) + sub * 8;
        kr[u] = ptx::ld_nc_v4(k_cache + off);
        vr[u] = ptx::ld_nc_v4(v_cache + off);
      }
    }
#pragma unroll
    for (int u = 0; u < UN; ++u) {
      const int pos = p0 + u;
      if (pos >= s_end) break;
      float kf[8], vf[8];
      unpack8<T>(kr[u], kf);
      unpack8<T>(vr[u], vf);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s = fmaf(qf[g][i], kf[i], s);
        // reduce over the 16 lanes of this half-warp
        // the two half-warps walk different key positions and may leave the loop at different times:
        // shuffle only inside this half-warp
        s += __shfl_xor_sync(half_mask, s, 8);
        s += __shfl_xor_sync(half_mask, s, 4);
        s += __shfl_xor_sync(half_mask, s, 2);
        s += __shfl_xor_sync(half_mask, s, 1);
        const float mn = fmaxf(m[g], s);
        const float corr = exp2f(m[g] - mn);
        const float pexp = exp2f(s - mn);
        l[g] = l[g] * corr + pexp;
#pragma unroll
        for (int i = 0; i < 8; ++i) o[g][i] = fmaf(pexp, vf[i], o[g][i] * corr);
        m[g] = mn;
      }
    }
  }
  // merge the 8 position groups through shared memory
  __shared__ float sh_m[8][G], sh_l[8][G];
  __shared__ float sh_o[8][G][D];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    if (sub == 0) {
      sh_m[grp][g] = m[g];
      sh_l[grp][g] = l[g];
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) sh_o[grp][g][sub * 8 + i] = o[g][i];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
    const int g = idx / D, d = idx % D;
    float mm = -CUDART_INF_F;
#pragma unroll
    for (int r = 0; r < 8; ++r) mm = fmaxf(mm, sh_m[r][g]);
    float ll = 0.f, acc = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const float w = sh_m[r][g] == -CUDART_INF_F ? 0.f : exp2f(sh_m[r][g] - mm);
      ll += sh_l[r][g] * w;
      acc += sh_o[r][g][d] * w;
    }
    float* dst = partial + ((size_t(b) * Hq + h0 + g) * splits + split) * (D + 1);
    dst[d] = ll > 0.f ? acc / ll : 0.f;
    if (d == 0) dst[D] = ll > 0.f ? (mm + log2f(ll)) * 0.6931471805599453f : -CUDART_INF_F;  // natural-log lse
  }
}

// Combine `splits` partials (out | lse) per (b, h): used for the intra-rank
// merge and — with splits = world size on the all-gathered buffer — for the
// inter-rank merge of distributed flash-decode (reference K51-K53).
// partial: [B, H, splits, 129] with strides (sb, sh, ss) in floats. out_bf: [B, H, 128] 16-bit (optional),
// out_f32: [B, H, 129] fp32 out | lse (optional, feeds the next level).
template <typename T>
__global__ void __launch_bounds__(128) flash_decode_combine_kernel(const float* __restrict__ partial, T* out16,
                                                                   float* out_f32, int H, int splits, long long sb,
                                                                   long long sh, long long ss) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  constexpr int D = 128;
  const int h = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
  const float* base = partial + size_t(b) * sb + size_t(h) * sh;
  float mx = -CUDART_INF_F;
  for (int s = 0; s < splits; ++s) mx = fmaxf(mx, base[size_t(s) * ss + D]);
  float den = 0.f, acc = 0.f;
  for (int s = 0; s < splits; ++s) {
    const float lse = base[size_t(s) * ss + D];
    const float w = lse == -CUDART_INF_F ? 0.f : __expf(lse - mx);
    den += w;
    acc += w * base[size_t(s) * ss + d];
  }
  const float r = den > 0.f ? acc / den : 0.f;
  if (out16) out16[(size_t(b) * H + h) * D + d] = Cvt<T>::one(r);
  if (out_f32) {
    float* o = out_f32 + (size_t(b) * H + h) * (D + 1);
    o[d] = r;
    if (d == 0) o[D] = den > 0.f ? mx + __logf(den) : -CUDART_INF_F;
  }
}

// ------------------------------------------------------------------ embedding gather, argmax
template <typename T>
__global__ void __launch_bounds__(256) embedding_kernel(const long long* __restrict__ ids, const T* __restrict__ table,
                                                        T* __restrict__ out, int hidden) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  const size_t t = blockIdx.x;
  const int4* src = reinterpret_cast<const int4*>(table + size_t(ids[t]) * hidden);
  int4* dst = reinterpret_cast<int4*>(out + t * hidden);
  for (int i = threadIdx.x; i < hidden / 8; i += blockDim.x) dst[i] = src[i];
}

template <typename T>
__global__ void __launch_bounds__(1024) argmax_kernel(const T* __restrict__ logits, long long* __restrict__ out, int vocab,
                                                      long long ld) {
  ptx::gdc_wait();  // programmatic dependent launch: the inputs come from the previous kernel in the stream
  ptx::gdc_launch_dependents();
  __shared__ float sv[32];
  __shared__ int si[32];
  const T* row = logits + size_t(blockIdx.x) * ld;
  float best = -CUDART_INF_F;
  int bi = 0;
  for (int i = threadIdx.x; i < vocab; i += blockDim.x) {
    const float v = Cvt<T>::f(row[i]);
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
  if ((threadIdx.x & 31) == 0) {
    sv[threadIdx.x >> 5] = best;
    si[threadIdx.x >> 5] = bi;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = blockDim.x >> 5;
    best = threadIdx.x < nw ? sv[threadIdx.x] : -CUDART_INF_F;
    bi = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) {
        best = ov;
        bi = oi;
      }
    }
    if (threadIdx.x == 0) out[blockIdx.x] = bi;
  }
}

}  // namespace tdist

namespace tdist {
struct FlashDecodeArgs {
  const void *q, *k, *v;
  const int* kv_lens;
  float* partial;
  void* out16;
  float* out_lse;
  int B, Hq, Hkv, splits;
  float sm_scale;
  long long sb, ss, sh;
  const int* block_table;
  int page_size, max_pages;
};
// Launch with the programmatic-dependent-launch attribute (off with TDIST_PDL=0): the decode path is a chain of ~10 small
// kernels per layer, each of which starts with gdc_wait(), so the launch + prologue of kernel i+1 overlaps the tail of
// kernel i (same effect inside a captured CUDA graph: the edges become programmatic).
struct PdlLaunch {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  PdlLaunch(dim3 grid, dim3 block, size_t smem, cudaStream_t s) {
    static const bool on = [] {
      const char* e = getenv("TDIST_PDL");
      return e == nullptr || atoi(e) != 0;
    }();
    cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = on ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  template <typename... KArgs, typename... Args>
  void operator()(void (*kern)(KArgs...), Args&&... args) {
    cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
  }
};

template <typename T, int G>
static void launch_fd_pass(const FlashDecodeArgs& a, int g_total, int g_off, cudaStream_t s) {
  PdlLaunch(dim3(a.splits, a.Hkv, a.B), 128, 0, s)(flash_decode_split_kernel<T, G>, 
      static_cast<const T*>(a.q), static_cast<const T*>(a.k), static_cast<const T*>(a.v), a.kv_lens, a.partial, a.Hq,
      a.Hkv, a.splits, a.sm_scale, a.sb, a.ss, a.sh, a.block_table, a.page_size, a.max_pages, g_total, g_off);
}
template <typename T>
static void launch_flash_decode(const FlashDecodeArgs& a, cudaStream_t s) {
  const int g = a.Hq / a.Hkv;
  int done = 0;
  while (done < g) {  // head groups larger than 8 take several passes
    const int rem = g - done;
    if (rem >= 8) {
      launch_fd_pass<T, 8>(a, g, done, s);
      done += 8;
    } else if (rem >= 4) {
      launch_fd_pass<T, 4>(a, g, done, s);
      done += 4;
    } else if (rem >= 2) {
      launch_fd_pass<T, 2>(a, g, done, s);
      done += 2;
    } else {
      launch_fd_pass<T, 1>(a, g, done, s);
      done += 1;
    }
  }
  if (a.out16 != nullptr || a.out_lse != nullptr) {
    PdlLaunch(dim3(a.Hq, a.B), 128, 0, s)(flash_decode_combine_kernel<T>, a.partial, static_cast<T*>(a.out16), a.out_lse, a.Hq,
                                                                   a.splits, (long long)a.Hq * a.splits * 129,
                                                                   (long long)a.splits * 129, 129);
  }
}
}  // namespace tdist

using namespace tdist;

#define DISPATCH_16(dtype, ...)                                   \
  if (dtype == TDIST_BF16) {                                      \
    using T = __nv_bfloat16;                                      \
    __VA_ARGS__;                                                  \
  } else if (dtype == TDIST_F16) {                                \
    using T = __half;                                             \
    __VA_ARGS__;                                                  \
  } else {                                                        \
    set_error("dtype must be bf16 or fp16");                      \
    return -1;                                                    \
  }

extern "C" {

int tdist_gemv(int dtype, const void* x, const void* w, const void* bias, void* y, int M, int N, int K, long long ldx,
               long long ldw, long long ldy, void* stream) {
  if (K % 8 != 0 || ldx % 8 != 0 || ldw % 8 != 0 || M < 1 || M > 16) {
    set_error("gemv: need 1 <= M <= 16 and K, ldx, ldw multiples of 8");
    return -1;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int warps = 8;
  int grid = (N + 2 * warps - 1) / (2 * warps);
  if (grid > 148 * 8) grid = 148 * 8;
  DISPATCH_16(dtype, {
    const T* xp = static_cast<const T*>(x);
    const T* wp = static_cast<const T*>(w);
    const T* bp = static_cast<const T*>(bias);
    T* yp = static_cast<T*>(y);
    if (M <= 1)
      PdlLaunch(grid, warps * 32, 0, s)(gemv_kernel<T, 1>, xp, wp, bp, yp, M, N, K, ldx, ldw, ldy);
    else if (M <= 2)
      PdlLaunch(grid, warps * 32, 0, s)(gemv_kernel<T, 2>, xp, wp, bp, yp, M, N, K, ldx, ldw, ldy);
    else if (M <= 4)
      PdlLaunch(grid, warps * 32, 0, s)(gemv_kernel<T, 4>, xp, wp, bp, yp, M, N, K, ldx, ldw, ldy);
    else if (M <= 8)
      PdlLaunch(grid, warps * 32, 0, s)(gemv_kernel<T, 8>, xp, wp, bp, yp, M, N, K, ldx, ldw, ldy);
    else
      PdlLaunch(grid, warps * 32, 0, s)(gemv_kernel<T, 16>, xp, wp, bp, yp, M, N, K, ldx, ldw, ldy);
  });
  return check_launch("gemv");
}

int tdist_rmsnorm(int dtype, const void* x, void* residual, const void* weight, void* y, int rows, int hidden, float eps,
                  long long ldx, long long ldy, void* stream) {
  if (hidden % 8 != 0 || ldx % 8 != 0 || ldy % 8 != 0) {
    set_error("rmsnorm: hidden and strides must be multiples of 8");
    return -1;
  }
  const size_t smem = (size_t(hidden) + 32) * sizeof(float);
  if (smem > 200 * 1024) {
    set_error("rmsnorm: hidden too large");
    return -1;
  }
  DISPATCH_16(dtype, {
    static bool attr = false;
    if (!attr) {
      cudaFuncSetAttribute(rmsnorm_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      attr = true;
    }
    PdlLaunch(rows, 256, smem, static_cast<cudaStream_t>(stream))(rmsnorm_kernel<T>, 
        static_cast<const T*>(x), static_cast<T*>(residual), static_cast<const T*>(weight), static_cast<T*>(y), hidden,
        eps, ldx, ldy);
  });
  return check_launch("rmsnorm");
}

int tdist_silu_mul(int dtype, const void* gate_up, void* out, long long tokens, int inter, void* stream) {
  if (inter % 8 != 0) {
    set_error("silu_mul: intermediate size must be a multiple of 8");
    return -1;
  }
  const long long total = tokens * (inter / 8);
  long long grid = (total + 255) / 256;
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  DISPATCH_16(dtype, {
    PdlLaunch(int(grid), 256, 0, static_cast<cudaStream_t>(stream))(silu_mul_kernel<T>, static_cast<const T*>(gate_up),
                                                                                static_cast<T*>(out), inter, total);
  });
  return check_launch("silu_mul");
}

int tdist_add(int dtype, const void* a, const void* b, void* out, long long numel, void* stream) {
  if (numel % 8 != 0) {
    set_error("add: numel must be a multiple of 8");
    return -1;
  }
  const long long nv = numel / 8;
  long long grid = (nv + 255) / 256;
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  DISPATCH_16(dtype, {
    PdlLaunch(int(grid), 256, 0, static_cast<cudaStream_t>(stream))(add_kernel<T>, static_cast<const T*>(a),
                                                                           static_cast<const T*>(b), static_cast<T*>(out), nv);
  });
  return check_launch("add");
}

int tdist_qk_norm_rope_kv(int dtype, const void* qkv, const void* q_norm_w, const void* k_norm_w, const float* cos_sin,
                          const int* positions, const int* batch_idx, void* q_out, void* k_cache, void* v_cache,
                          int tokens, int Hq, int Hkv, int head_dim, float eps, long long kv_batch_stride,
                          long long kv_seq_stride, long long kv_head_stride, const int* block_table, int page_size,
                          int max_pages, void* stream) {
  if (head_dim != 128) {
    set_error("qk_norm_rope_kv: head_dim must be 128");
    return -1;
  }
  const long long warps = (long long)tokens * (Hq + 2 * Hkv);
  const int grid = int((warps * 32 + 255) / 256);
  if (grid == 0) return 0;
  DISPATCH_16(dtype, {
    PdlLaunch(grid, 256, 0, static_cast<cudaStream_t>(stream))(qk_norm_rope_kv_kernel<T>, 
        static_cast<const T*>(qkv), static_cast<const T*>(q_norm_w), static_cast<const T*>(k_norm_w), cos_sin, positions,
        batch_idx, static_cast<T*>(q_out), static_cast<T*>(k_cache), static_cast<T*>(v_cache), tokens, Hq, Hkv, eps,
        kv_batch_stride, kv_seq_stride, kv_head_stride, block_table, page_size, max_pages);
  });
  return check_launch("qk_norm_rope_kv");
}

int tdist_flash_decode(int dtype, const void* q, const void* k_cache, const void* v_cache, const int* kv_lens,
                       float* partial, void* out16, float* out_lse, int B, int Hq, int Hkv, int head_dim, int splits,
                       float sm_scale, long long kv_batch_stride, long long kv_seq_stride, long long kv_head_stride,
                       const int* block_table, int page_size, int max_pages, void* stream) {
  if (head_dim != 128 || Hq % Hkv != 0) {
    set_error("flash_decode: head_dim must be 128 and Hq a multiple of Hkv");
    return -1;
  }
  FlashDecodeArgs a{q, k_cache, v_cache, kv_lens, partial, out16, out_lse, B, Hq, Hkv, splits, sm_scale,
                    kv_batch_stride, kv_seq_stride, kv_head_stride, block_table, page_size, max_pages};
  DISPATCH_16(dtype, { launch_flash_decode<T>(a, static_cast<cudaStream_t>(stream)); });
  return check_launch("flash_decode");
}

int tdist_flash_decode_combine(int dtype, const float* partial, void* out16, float* out_lse, int B, int H, int splits,
                               long long stride_b, long long stride_h, long long stride_s, void* stream) {
  DISPATCH_16(dtype, {
    PdlLaunch(dim3(H, B), 128, 0, static_cast<cudaStream_t>(stream))(flash_decode_combine_kernel<T>, 
        partial, static_cast<T*>(out16), out_lse, H, splits, stride_b, stride_h, stride_s);
  });
  return check_launch("flash_decode_combine");
}

int tdist_embedding(int dtype, const long long* ids, const void* table, void* out, int tokens, int hidden, void* stream) {
  if (hidden % 8 != 0) {
    set_error("embedding: hidden must be a multiple of 8");
    return -1;
  }
  if (tokens == 0) return 0;
  DISPATCH_16(dtype, {
    PdlLaunch(tokens, 256, 0, static_cast<cudaStream_t>(stream))(embedding_kernel<T>, ids, static_cast<const T*>(table),
                                                                              static_cast<T*>(out), hidden);
  });
  return check_launch("embedding");
}

int tdist_argmax(int dtype, const void* logits, long long* out, int rows, int vocab, long long ld, void* stream) {
  if (rows == 0) return 0;
  DISPATCH_16(dtype, {
    PdlLaunch(rows, 1024, 0, static_cast<cudaStream_t>(stream))(argmax_kernel<T>, static_cast<const T*>(logits), out, vocab, ld);
  });
  return check_launch("argmax");
}

}  // extern "C"

// ------------------------------------------------------------------ flash decode without any allocation
//
// tdist_flash_decode() needs a caller-provided fp32 scratch buffer for the per-split partial results. Some callers
// cannot provide one: a serving loop that has handed all device memory to its own pool, plugins that are loaded into a
// process whose allocator they must not touch, and C callers that have no allocator wired up at all. For them the scratch below lives inside the library image: the driver
// reserves it once per GPU when the module is loaded and keeps it until the process exits, so this entry point never
// calls cudaMalloc / cudaFree (nor any torch or library routine that could).
//
// Capacity: kNoAllocMaxRows partial rows of 129 floats (one row per batch x query head x kv split), i.e. ~2.0 MiB per
// GPU. Larger problems are rejected with an error; use tdist_flash_decode() with your own buffer for those.
//
// Two things to know when calling it:
//  * The first call on each GPU looks up where that GPU's copy of the scratch lives (cudaGetSymbolAddress, with the
//    calling thread's current device). That lookup is a host-side query which may load the module, so the first call
//    on a device has to come before any stream capture is started; later calls on that device only launch kernels.
//  * There is ONE scratch per GPU, shared by every call and left holding the previous call's partials. Calls issued to
//    the same stream are ordered and safe. Two calls that can run at the same time on one GPU (different streams,
//    different host threads) would overwrite each other's partials: serialise them or use tdist_flash_decode().
namespace tdist {
constexpr int kNoAllocMaxRows = 4096;
static __device__ __align__(256) float g_flash_decode_scratch[kNoAllocMaxRows * 129];

// In host code `g_flash_decode_scratch` is only a handle for the symbol, not an address the GPU can use: the device
// address has to be looked up per device (every GPU that loads the library has its own copy of the array).
constexpr int kMaxDevices = 64;
static float* g_scratch_dev_ptr[kMaxDevices] = {nullptr};

static int no_alloc_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error(std::string("flash_decode_noalloc: cannot identify the current device: ") + cudaGetErrorString(e));
    return -1;
  }
  if (g_scratch_dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_flash_decode_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error(std::string("flash_decode_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
      return -1;
    }
    g_scratch_dev_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_dev_ptr[dev];
  return 0;
}
}  // namespace tdist

extern "C" int tdist_flash_decode_noalloc(int dtype, const void* q, const void* k_cache, const void* v_cache,
                                          const int* kv_lens, void* out16, int B, int Hq, int Hkv, int head_dim, int splits,
                                          float sm_scale, long long kv_batch_stride, long long kv_seq_stride,
                                          long long kv_head_stride, const int* block_table, int page_size, int max_pages,
                                          void* stream) {
  if (head_dim != 128 || Hkv <= 0 || Hq % Hkv != 0) {
    set_error("flash_decode_noalloc: head_dim must be 128 and Hq a multiple of Hkv");
    return -1;
  }
  if (B <= 0 || splits <= 0 || out16 == nullptr) {
    set_error("flash_decode_noalloc: B and splits must be positive and out16 must be given");
    return -1;
  }
  const long long rows = (long long)B * Hq * splits;
  if (rows > kNoAllocMaxRows) {
    set_error("flash_decode_noalloc: B * Hq * splits = " + std::to_string(rows) + " exceeds the built-in scratch of " +
              std::to_string(kNoAllocMaxRows) + " rows; use tdist_flash_decode() with a caller-provided buffer");
    return -1;
  }
  float* scratch = nullptr;
  if (no_alloc_scratch(&scratch) != 0) return -1;
  // Nothing has to be cleared: the split kernel writes every (batch, head, split) row it is launched for, including
  // the "empty split" marker, and the combine kernel reads exactly those rows.
  FlashDecodeArgs a{q, k_cache, v_cache, kv_lens, scratch, out16, nullptr, B, Hq, Hkv, splits, sm_scale,
                    kv_batch_stride, kv_seq_stride, kv_head_stride, block_table, page_size, max_pages};
  DISPATCH_16(dtype, { launch_flash_decode<T>(a, static_cast<cudaStream_t>(stream)); });
  return check_launch("flash_decode_noalloc");
}
