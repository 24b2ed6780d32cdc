// primitives_selftest.cu — one small kernel that uses every NVSHMEM-style device primitive of csrc/device/symm.cuh
// (my_pe / n_pes / symm_at / int_p / putmem / getmem / putmem_signal / signal_op / signal_wait_until / broadcast /
// fcollect / fence / quiet / team_translate_pe), so that the whole header is instantiated at build time and can be run
// as a test (triton_dist/test/nvidia/test_device_primitives.py). No counterpart in the reference: there the same
// coverage comes from test_nvshmem_api.py / test_ring_put.py calling NVSHMEM through Triton.
#include <cuda_runtime.h>

#include "../../include/tdist.h"
#include "../device/symm.cuh"
#include "common.cuh"

namespace tdist {

// Layout of the symmetric scratch `buf` in 16-byte units (int4): [0, W) ring slots, [W, 2W) fcollect destination,
// [2W, 3W) broadcast buffer, [3W] result, [3W + 1] my source value. flags: [0, W) put signals, [W] broadcast signal,
// [W + 1, 2W + 1) fcollect signals. One CTA of 128 threads per rank; `epoch` grows with every call (nothing is reset).
__global__ void __launch_bounds__(128) primitives_selftest_kernel(SymmCtx c, int4* buf, uint32_t* flags, uint32_t epoch) {
  const int W = n_pes(c), me = my_pe(c), nxt = (me + 1) % W, prv = (me + W - 1) % W;
  const int e = int(epoch);
  int4* src = buf + 3 * W + 1;
  if (threadIdx.x == 0) *src = make_int4(me * 10 + e, me, e, 7);
  __syncthreads();
  // ring put with signal: my value into slot `me` of the next rank, then wait for the previous rank's signal
  putmem_signal_block(c, buf + me, src, sizeof(int4), flags + me, epoch, SIGNAL_SET, nxt);
  if (threadIdx.x == 0) signal_wait_until(flags + prv, CMP_GE, epoch);
  __syncthreads();
  // fcollect: everybody's value at [W + rank] everywhere, completion through signal_op + signal_wait_until
  fcollect_block(c, buf + W, src, sizeof(int4));
  __syncthreads();
  if (threadIdx.x == 0) {
    fence();
    for (int p = 0; p < W; ++p) signal_op(c, flags + W + 1 + me, epoch, SIGNAL_SET, p);
    for (int p = 0; p < W; ++p) signal_wait_until(flags + W + 1 + p, CMP_GE, epoch);
  }
  __syncthreads();
  // broadcast from rank 0 (NVLS store stream when the heap has a multicast mapping)
  if (me == 0 && threadIdx.x < W) buf[2 * W + threadIdx.x] = make_int4(1000 + e + int(threadIdx.x), 0, 0, 0);
  __syncthreads();
  broadcast_block(c, buf + 2 * W, size_t(W) * sizeof(int4), 0);
  __syncthreads();
  if (threadIdx.x == 0) {
    quiet();
    if (me == 0)
      for (int p = 0; p < W; ++p) signal_op(c, flags + W, epoch, SIGNAL_SET, p);
    signal_wait_until(flags + W, CMP_GE, epoch);
    int ok = ptx::ld_v4(buf + prv).x == prv * 10 + e;
    for (int p = 0; p < W; ++p) ok = ok && ptx::ld_v4(buf + W + p).x == p * 10 + e;
    for (int p = 0; p < W; ++p) ok = ok && ptx::ld_v4(buf + 2 * W + p).x == 1000 + e + p;
    ok = ok && team_translate_pe(0, 1, me, 0, 2, (W + 1) / 2) == (me % 2 == 0 ? me / 2 : -1);
    ok = ok && ptx::ld_v4(remote_ptr(c, buf + W + nxt, nxt)).x == nxt * 10 + e;  // load through the peer mapping = get
    int_p(c, reinterpret_cast<int*>(buf + 3 * W), ok ? e : -1, me);
  }
  // instantiation of the warp-level forms (zero bytes: no data movement)
  if (threadIdx.x < 32) {
    getmem_warp(c, buf + 3 * W + 1, buf + W + nxt, 0, nxt);
    putmem_nbi_warp(c, buf + me, src, 0, nxt);
    putmem_signal_warp(c, buf + me, src, 0, flags + me, epoch, SIGNAL_SET, me);
  }
  putmem_nbi_block(c, buf + me, src, 0, nxt);
  getmem_block(c, buf + 3 * W + 1, buf + W + me, 0, me);
}

}  // namespace tdist

extern "C" int tdist_primitives_selftest(const tdist_symm_ctx* ctx, void* buf, uint32_t* flags, uint32_t epoch, void* stream) {
  tdist::SymmCtx c;
  for (int i = 0; i < TDIST_MAX_RANKS; ++i) c.base[i] = static_cast<char*>(ctx->base[i]);
  c.mc_base = static_cast<char*>(ctx->mc_base);
  c.bytes = ctx->bytes;
  c.rank = ctx->rank;
  c.world = ctx->world;
  tdist::primitives_selftest_kernel<<<1, 128, 0, static_cast<cudaStream_t>(stream)>>>(c, static_cast<int4*>(buf), flags, epoch);
  return tdist::check_launch("primitives_selftest");
}
