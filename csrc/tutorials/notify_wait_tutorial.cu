// notify_wait_tutorial.cu — tutorial 1: a producer / consumer queue between neighbouring GPUs, built from nothing but
// symm_at + notify + wait (csrc/device/symm.cuh). Counterpart of the reference's tutorials/01-distributed-notify-wait.py,
// whose Triton kernel moves blocks through a small queue with dl.wait / dl.consume_token / dl.notify.
//
// One launch per rank, two CTAs: CTA 0 PRODUCES items for the next rank, CTA 1 CONSUMES the items the previous rank
// produces for this one. (Both roles live in ONE launch on purpose: a kernel that spins on a flag must never depend on a
// LATER launch of the same GPU getting an SM.) The queue of rank r lives in r's symmetric heap:
//
//     queue  [slots][item_vecs] int4      payload slots, written by the producer on rank r-1, read by r's consumer
//     full   [slots] u32                  "slot holds item with this sequence number"; written by r-1, polled by r
//     empty  [slots] u32                  lives on the PRODUCER's rank: "the consumer has drained this slot up to ..."
//
// Sequence numbers instead of booleans: item i uses slot i % slots in round i / slots, and the value that marks it is
// `base + round + 1`, where `base` grows from call to call. Nothing is ever reset, so a rank that is already in the next
// call cannot be confused with the previous one, and the kernel needs no barrier at its start or end.
#include "../kernels/common.cuh"

namespace tdist {
namespace tut {

constexpr int kQueueThreads = 256;

__global__ void __launch_bounds__(kQueueThreads) tut_queue_kernel(SymmCtx ctx, const int4* in, int4* out, int4* queue,
                                                                  uint32_t* full, uint32_t* empty, int n_items, int item_vecs,
                                                                  int slots, uint32_t base) {
  const int nxt = (ctx.rank + 1) % ctx.world;
  const int prv = (ctx.rank + ctx.world - 1) % ctx.world;
  const int rounds = (n_items + slots - 1) / slots;
  if (blockIdx.x == 0) {
    // ------------------------------------------------------------------ producer: my `in` -> queue of rank `nxt`
    int4* peer_queue = symm_at(ctx, queue, nxt);  // the same offset in nxt's heap: plain address arithmetic
    for (int i = 0; i < n_items; ++i) {
      const int slot = i % slots, round = i / slots;
      // (1) wait until the consumer has drained what this slot held before: `empty` is on MY rank, written by nxt.
      //     One thread polls (acquire), the barrier releases the rest of the CTA.
      if (threadIdx.x == 0) wait_ge_sys(empty + slot, base + round);
      __syncthreads();
      // (2) payload: ordinary 16-byte stores through the peer mapping = a one-sided put over NVLink
      for (int v = threadIdx.x; v < item_vecs; v += kQueueThreads)
        ptx::st_v4(peer_queue + size_t(slot) * item_vecs + v, ptx::ld_nc_v4(in + size_t(i) * item_vecs + v));
      // (3) publish: all threads' stores are issued (barrier), ONE thread makes them visible system-wide (fence) and only
      //     then writes the flag on the consumer's rank. This is the whole "notify" contract of the library.
      __syncthreads();
      if (threadIdx.x == 0) {
        fence();
        notify(ctx, full + slot, nxt, base + round + 1);
      }
    }
  } else {
    // ------------------------------------------------------------------ consumer: my queue -> my `out`
    for (int i = 0; i < n_items; ++i) {
      const int slot = i % slots, round = i / slots;
      if (threadIdx.x == 0) wait_ge_sys(full + slot, base + round + 1);  // acquire: the payload is visible after this
      __syncthreads();
      for (int v = threadIdx.x; v < item_vecs; v += kQueueThreads)
        out[size_t(i) * item_vecs + v] = ptx::ld_v4(queue + size_t(slot) * item_vecs + v);
      // hand the slot back to the producer on `prv`: my loads of the slot must be complete before it may be overwritten
      __syncthreads();
      if (threadIdx.x == 0) {
        fence();
        notify(ctx, empty + slot, prv, base + round + 1);
      }
    }
    // close the call: every slot is drained up to the NEXT call's base, whatever round it was last used in
    if (threadIdx.x < slots) notify(ctx, empty + threadIdx.x, prv, base + rounds);
  }
}

}  // namespace tut
}  // namespace tdist

using namespace tdist;

extern "C" {
/* tutorials/01: rank r streams n_items items of item_vecs 16-byte vectors from `in` into the `out` of rank r+1 through a
 * queue of `slots` slots in r+1's heap. queue [slots][item_vecs] int4, full / empty [slots] u32: symmetric, zeroed once.
 * `base` = sum of ceil(n_items / slots) over all earlier calls on these buffers (0 for the first call). */
int tdist_tutorial_queue(const tdist_symm_ctx* ctx, const void* in, void* out, void* queue, uint32_t* full, uint32_t* empty,
                         int n_items, int item_vecs, int slots, uint32_t base, void* stream) {
  if (n_items < 1 || item_vecs < 1 || slots < 1 || slots > tut::kQueueThreads) {
    set_error("tutorial_queue: n_items, item_vecs >= 1, 1 <= slots <= 256");
    return -1;
  }
  tut::tut_queue_kernel<<<2, tut::kQueueThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      to_ctx(ctx), static_cast<const int4*>(in), static_cast<int4*>(out), static_cast<int4*>(queue), full, empty, n_items,
      item_vecs, slots, base);
  return check_launch("tutorial_queue");
}
}
