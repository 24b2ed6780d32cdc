"""Tutorial 1 — symmetric memory, notify and wait: a producer / consumer queue between GPUs.
(Counterpart of the reference's tutorials/01-distributed-notify-wait.py, which teaches ``dl.wait`` /
``dl.consume_token`` / ``dl.notify`` with a Triton kernel that streams blocks through a small queue. Here the same
three ideas are CUDA device functions of csrc/device/symm.cuh and the queue kernel is
csrc/tutorials/notify_wait_tutorial.cu.)

    bash scripts/launch.sh tutorials/01-distributed-notify-wait.py           # all GPUs of the node
    bash scripts/launch.sh --nproc_per_node=2 tutorials/01-distributed-notify-wait.py

You will learn

  * how the distributed runtime starts and what a SYMMETRIC tensor is;
  * the three device primitives everything else in this library is made of: ``symm_at``, ``notify``, ``wait``;
  * why flags carry sequence numbers instead of booleans;
  * the same signalling from the HOST side (stream memory operations), which is what copy-engine producers use.

STEP 1 — the runtime. ``initialize_distributed()`` starts torch.distributed (one process per GPU; NCCL is only glue and
baseline), then creates the symmetric heap (csrc/runtime/symm_heap.cc): every rank ``cuMemCreate``s the same number of
bytes, the ranks exchange file descriptors, and every rank maps ALL W allocations into one virtual address range, rank p
at ``base + p * heap_bytes``. ``symm_tensor(shape, dtype)`` carves the same offset out of every rank's heap (all ranks
must allocate in the same order — it is a collective in the sense of NVSHMEM's ``nvshmem_malloc``). Consequence:

        T* symm_at(ctx, T* local, int peer)  =  local + (peer - ctx.rank) * heap_bytes          // no table, no lookup

A kernel on rank r can load and store rank p's copy of a symmetric tensor directly; NVLink and the switch do the rest.
From Python the same alias is ``symm_peer_view(t, p)``: a torch tensor over the peer's memory.

STEP 2 — signals. A flag is a 32-bit word in the heap. The contract (csrc/device/symm.cuh):

        producer                                         consumer
        --------                                         --------
        payload stores (any number of threads)           one thread:  wait_ge_sys(flag, value)    // ld.acquire.sys in a loop
        __syncthreads()                                  __syncthreads()
        one thread:  fence()                             loads of the payload
                     notify(ctx, flag, peer, value)      // st.release.sys on the peer's copy of `flag`

The barrier collects the CTA's stores, the fence makes them visible system-wide, and only then the flag is written;
on the other side the acquire load orders the payload loads after the flag. (The reference lowers ``dl.notify`` to
``membar.sys`` + ``st.relaxed.sys`` by thread 0 and ``dl.wait`` to an acquire spin by the first lanes of each warp followed
by a warp barrier — DistributedOpToLLVM.cpp:179-336; ``consume_token`` only pins that order for the Triton scheduler and
has no counterpart in CUDA, where program order already holds.)

STEP 3 — the queue kernel. Rank r produces ``n_items`` items for rank r+1 through a queue of ``slots`` slots that lives
in r+1's heap. Per slot there are two flags: ``full`` (on the consumer's rank, written by the producer) and ``empty``
(on the producer's rank, written by the consumer). CTA 0 of the launch is the producer, CTA 1 the consumer of the
queue that the previous rank fills:

        producer, item i (slot s = i % slots, round q = i / slots)        consumer, item i
        wait   empty[s] >= base + q          (slot drained)               wait   full[s] >= base + q + 1
        put    item -> queue[s] on rank r+1                               copy   queue[s] -> out[i]
        notify full[s] on r+1 := base + q + 1                             notify empty[s] on r-1 := base + q + 1

STEP 4 — sequence numbers. ``base`` grows by ceil(n_items / slots) with every call and the flags are NEVER reset. A
boolean flag would need "everybody has seen it, now clear it, now everybody has seen the clear" — two more barriers per
call. With monotone values a rank that is already in call n+1 cannot be mistaken for call n, so calls follow each other
with no barrier in between and the launch can be captured in a CUDA graph. Every fused op of this library uses the same
trick (``PhaseState`` in symm.cuh keeps the counter on the device). This kernel, unmodified, also runs on the CPU with
1-5 simulated ranks under ThreadSanitizer (csrc/sim, tests/test_cpu_protocol_sim.py): no payload access without a
release / acquire chain through the flags, no lost wake-up, with more and with fewer items than slots.

STEP 5 — the host side of the same protocol: ``set_signal`` / ``wait_eq`` enqueue a flag write / a flag wait on the
current stream (one-thread kernels by default, so that they can be captured in CUDA graphs; ``use_memops=True`` selects
``cuStreamWriteValue32`` / ``cuStreamWaitValue32``): a copy into a peer view followed by ``set_signal`` is a
put-with-signal in stream order, and ``wait_eq`` blocks the STREAM (not the CPU) until the flag arrives.
"""
import ctypes
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist

from triton_dist import _lib
from triton_dist.kernels.common_ops import CMP_GE, set_signal, wait_eq
from triton_dist.runtime import get_heap
from triton_dist.utils import (dist_print, finalize_distributed, initialize_distributed, nvshmem_barrier_all_on_stream,
                               symm_peer_view, symm_tensor)

# ---- step 1: runtime + symmetric tensors
initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
if not torch.cuda.is_available():
    dist_print("this tutorial needs GPUs: on a CPU-only machine the package runs its ops through torch.distributed (gloo) "
               "and has no device primitives to show")
    finalize_distributed()
    raise SystemExit(0)
heap = get_heap()
nxt, prv = (rank + 1) % world, (rank - 1) % world
n_items, item_elems, slots = 64, 4096, 4            # 64 items of 16 KiB through a 4-slot queue
item_vecs = item_elems * 4 // 16
queue = symm_tensor((slots, item_elems), torch.float32, zero=False)
full = symm_tensor((slots,), torch.int32)             # zero-filled: "nothing produced yet"
empty = symm_tensor((slots,), torch.int32)            # zero-filled: "every slot drained up to sequence 0"
nvshmem_barrier_all_on_stream()                       # every rank has zeroed its flags before anybody signals
dist_print(f"heap mode {heap.mode} (1 = VMM, 2 = cudaIpc), multicast {bool(heap.multicast)}; my queue is at offset "
           f"{heap.offset_of(queue):#x} of EVERY rank's heap")

# ---- steps 2-4: the queue kernel, called several times back to back without any barrier or reset
lib = _lib.lib()
lib.tdist_tutorial_queue.restype = ctypes.c_int
lib.tdist_tutorial_queue.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int] * 3 + [ctypes.c_uint32, ctypes.c_void_p]
out = torch.empty(n_items, item_elems, device="cuda")
base = 0
for call in range(3):
    src = torch.full((n_items, item_elems), float(100 * call + rank), device="cuda")
    src += torch.arange(n_items, device="cuda", dtype=torch.float32)[:, None] / 1024   # item i carries i in its fraction
    _lib.check(lib.tdist_tutorial_queue(ctypes.byref(heap.ctx), src.data_ptr(), out.data_ptr(), queue.data_ptr(),
                                        full.data_ptr(), empty.data_ptr(), n_items, item_vecs, slots, base, _lib.stream_ptr()),
               "tutorial_queue")
    base += (n_items + slots - 1) // slots            # step 4: the next call starts where this one's sequence numbers ended
    expect = torch.full((n_items, item_elems), float(100 * call + prv), device="cuda")
    expect += torch.arange(n_items, device="cuda", dtype=torch.float32)[:, None] / 1024
    ok = torch.equal(out, expect)                     # same stream: the comparison runs after the kernel
    dist_print(f"rank {rank} call {call}: {n_items} items from rank {prv} through {slots} slots: {'ok' if ok else 'WRONG'}",
               allowed_ranks="all")
    assert ok

# ---- step 5: the same put + signal from the host, no kernel of ours involved
data = symm_tensor((1024,), torch.float32)
flag = symm_tensor((world,), torch.int32)
nvshmem_barrier_all_on_stream()
symm_peer_view(data, nxt).fill_(float(rank))                      # put: a fill kernel writing rank nxt's memory
set_signal(symm_peer_view(flag, nxt)[rank:rank + 1], 1)          # stream-ordered after the put
wait_eq(flag[prv:prv + 1], 1, cmp=CMP_GE)                        # my STREAM waits for my left neighbour's signal
got = data[0].item()                                             # .item() synchronises the stream
dist_print(f"rank {rank}: host-side put + signal delivered {got} from rank {prv}", allowed_ranks="all")
assert got == float(prv)
torch.cuda.synchronize()
nvshmem_barrier_all_on_stream()
torch.cuda.synchronize()
for t in (flag, data, empty, full, queue):
    heap.free(t)
finalize_distributed()
