"""Tutorial 2 — all-gather inside one NVSwitch node, four ways, from copy engines to LL packets.
(Counterpart of the reference's tutorials/02-intra-node-allgather.py, which gathers with copy-engine pulls / pushes on
side streams and signals the consumer with ``cuStreamWriteValue``. The kernels used here: ``allgather_kernel`` and
``allgather_ll_kernel`` in csrc/kernels/collectives.cu.)

    bash scripts/launch.sh tutorials/02-intra-node-allgather.py

The problem: rank r holds ``x_r [m, K]``; afterwards every rank holds ``[x_0; x_1; ...; x_{W-1}]``. It is the producer
half of AllGather + GEMM (tutorial 7), so what matters is not only the time of the whole gather but WHEN each shard
becomes usable and how the consumer learns about it.

STEP 0 — baseline: ``torch.distributed.all_gather_into_tensor`` (NCCL). One call, one completion event: nothing can
be consumed before everything has arrived.

STEP 1 — copy engines + stream signals (``cp_engine_producer_all_gather_intra_node``). Every rank owns a symmetric
``[W * m, K]`` buffer; ``bufs[p]`` is rank p's instance mapped into this process. On a side stream, for every source
rank: ``cudaMemcpyAsync`` of that rank's shard out of ITS buffer into mine (a pull over NVLink by a copy engine, no SM
involved), followed by ``cuStreamWriteValue32(signal[src], epoch)``. A consumer stream enqueues a wait on
``signal[src]`` before the work that needs shard ``src`` (``wait_eq``: by default a one-thread kernel, which can be
captured in a CUDA graph; ``use_memops=True`` selects ``cuStreamWaitValue32``): per-shard readiness without the host.
This is the reference's default producer; its cost is ~2-3 us of launch / DMA set-up per copy (W-1 copies), so it is the
right tool for LARGE shards when the SMs are busy with a GEMM.

STEP 2 — one push kernel over P2P stores (mode ``push2d`` / ``push3d``): each CTA owns a slice of the shard, loads it once
and stores it into the same offset of every peer's buffer (``symm_at``), 8 independent 16-byte loads in flight per thread.
A per-CTA cross-rank flag barrier (``barrier_all_block`` on that CTA's flag row) ends the kernel: CTA b of every rank
handles slice b, so "slice b has arrived from everybody" needs no grid-wide synchronisation. W-1 copies of the shard
leave the GPU.

STEP 3 — the same kernel over NVLS multicast (mode ``push_multimem``): the heap is also mapped through a multicast object
(``cuMulticastCreate`` / ``cuMulticastBindMem`` in csrc/runtime/symm_heap.cc), and ONE ``multimem.st`` to
``mc_ptr(ctx, buf)`` is replicated by the switch into all W copies. The shard leaves the GPU once: 1/(W-1) of the
store instructions and of the uplink traffic. Measured on 8 x B200 (profiles/README.md, link probe): 706 GB/s of ingress
per GPU with 8 CTAs against 287 GB/s for unicast stores from the same 8 CTAs.

STEP 4 — LL packets for small messages (mode ``push2d_ll``, ``push2d_ll_multimem``). Below ~100 KB the time is latency:
store, fence, flag store, and on the other side flag poll, then payload load — two NVLink round trips. An LL packet is
16 bytes {data0, flag, data1, flag}, written with ONE 16-byte store (a 16-byte aligned store is delivered whole), so the
payload is its own signal: the receiver spins on the packet until both flag words carry this call's sequence number.
One one-way trip, no fence, no separate flag; the price is 2x the bytes. The packet area has two halves used by
alternating calls (a fixed half-buffer apart, whatever the message size), so no barrier is needed between calls.

Which one when (8 x B200, us, NCCL in brackets; profiles/README.md R2-4 and round-1 section 4): 4 KB 7.8 (21.5),
64 KB 11.8 (20.8), 1 MB 13.1 (24.3), 64 MB 107 (139) = 547 GB/s.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist

from triton_dist.kernels.common_ops import CMP_GE, wait_eq
from triton_dist.kernels.nvidia import cp_engine_producer_all_gather_intra_node, create_fast_allgather_context, fast_allgather
from triton_dist.utils import (dist_print, finalize_distributed, initialize_distributed, nvshmem_barrier_all_on_stream,
                               nvshmem_create_tensors, perf_func_max_over_ranks, symm_tensor)

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
if not torch.cuda.is_available():
    dist_print("this tutorial needs GPUs (peer mappings, copy engines, multicast)")
    finalize_distributed()
    raise SystemExit(0)
m, K = 1024, 4096
torch.manual_seed(rank)
x = torch.randn(m, K, device="cuda").bfloat16()

# ---- step 0: the oracle (bit-exact: an all-gather moves bytes, it does not compute)
ref = torch.empty(world * m, K, device="cuda", dtype=torch.bfloat16)
dist.all_gather_into_tensor(ref, x)


def nccl_all_gather():
    dist.all_gather_into_tensor(ref, x)


# ---- step 1: copy engines + stream signals
bufs = nvshmem_create_tensors((world * m, K), torch.bfloat16, rank, world)   # bufs[p] = rank p's buffer, mapped here
sigs = nvshmem_create_tensors((world,), torch.int32, rank, world)            # sigs[rank][src] = "shard src has landed"
ag_stream = torch.cuda.Stream()
epoch = 0


def copy_engine_all_gather():
    """Producer on the side stream, per-shard waits on the current stream (where a consumer GEMM would run)."""
    global epoch
    epoch += 1                                          # monotone signal value: nothing is reset between calls
    nvshmem_barrier_all_on_stream()                     # no peer is still pulling the previous call's shard out of my slot
    bufs[rank][rank * m:(rank + 1) * m].copy_(x)        # my shard into my slot
    nvshmem_barrier_all_on_stream()                     # every rank's slot is filled before anybody pulls from it
    ag_stream.wait_stream(torch.cuda.current_stream())
    cp_engine_producer_all_gather_intra_node(rank, world, x, bufs, ag_stream, sigs, signal_target=epoch)
    for src in range(world):
        if src != rank:
            wait_eq(sigs[rank][src:src + 1], epoch, cmp=CMP_GE)   # the STREAM waits for the DMA's signal, not the host
            # ... a GEMM over rows [src * m, (src + 1) * m) would be enqueued here ...
    return bufs[rank]


got = copy_engine_all_gather()
ok_ce = torch.equal(got, ref)

# ---- steps 2-4: the kernels. A symmetric [W, shard] buffer, slot `rank` filled locally, all slots valid on return.
ctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=1 << 20)   # the LL packet area; the other modes need none
big = symm_tensor((world, m, K), torch.bfloat16)
small = symm_tensor((world, 512), torch.int32)         # 2 KB per rank: the LL regime
nvshmem_barrier_all_on_stream()                        # everybody's zero fill is done before the first push arrives
results = {}
for mode in ("pull", "push2d", "push_multimem"):
    big[rank].copy_(x)
    fast_allgather(big, ctx, mode=mode)
    results[mode] = torch.equal(big.view(world * m, K), ref)
    nvshmem_barrier_all_on_stream()                    # the next mode overwrites slots that peers may still be checking
for mode in ("push2d_ll", "push2d_ll_multimem"):
    small[rank].fill_(rank + 1)
    fast_allgather(small, ctx, mode=mode)
    results[mode] = all(bool((small[s] == s + 1).all()) for s in range(world))
    nvshmem_barrier_all_on_stream()
torch.cuda.synchronize()
dist_print(f"rank {rank}: copy engines {'ok' if ok_ce else 'WRONG'} | " +
           " | ".join(f"{k} {'ok' if v else 'WRONG'}" for k, v in results.items()), allowed_ranks="all")
assert ok_ce and all(results.values())

# ---- timings (device time, max over ranks). The ops synchronise across ranks themselves, so they can be called back to back.
big[rank].copy_(x)
_, t_nccl = perf_func_max_over_ranks(nccl_all_gather, 20, 5)
_, t_ce = perf_func_max_over_ranks(copy_engine_all_gather, 20, 5)
_, t_push = perf_func_max_over_ranks(lambda: fast_allgather(big, ctx, mode="push2d"), 20, 5)
_, t_mc = perf_func_max_over_ranks(lambda: fast_allgather(big, ctx, mode="push_multimem"), 20, 5)
_, t_ll = perf_func_max_over_ranks(lambda: fast_allgather(small, ctx, mode="push2d_ll"), 20, 5)
small_ref = torch.empty(world * 512, device="cuda", dtype=torch.int32)
_, t_ll_nccl = perf_func_max_over_ranks(lambda: dist.all_gather_into_tensor(small_ref, small[rank]), 20, 5)
mb = world * m * K * 2 / 2**20
dist_print(f"{mb:.0f} MiB gathered: NCCL {t_nccl * 1e3:.1f} us | copy engines + signals {t_ce * 1e3:.1f} us (includes its two barriers) | "
           f"P2P push {t_push * 1e3:.1f} us | multicast push {t_mc * 1e3:.1f} us")
dist_print(f"{world * 2} KiB gathered: NCCL {t_ll_nccl * 1e3:.1f} us | LL packets {t_ll * 1e3:.1f} us")
torch.cuda.synchronize()
ctx.finalize()
finalize_distributed()
