"""Tutorial 3 — the hierarchical ("inter-node") all-gather modes, and what is left of them on one NVSwitch node.
(Counterpart of the reference's tutorials/03-inter-node-allgather.py, which crosses nodes with NVSHMEM
``putmem_signal`` over InfiniBand and forwards inside the node over NVLink.)

    bash scripts/launch.sh tutorials/03-inter-node-allgather.py

WHY THE REFERENCE HAS A HIERARCHY. On a cluster of 8-GPU nodes a rank reaches 7 peers over NVLink (hundreds of GB/s)
and the rest over a NIC (50 GB/s, shared). A flat all-gather would send every shard W - 1 times through the NIC. The
reference therefore works in two (or three) levels:

    2-D   (node, local rank):   1. rank (n, l) sends its shard ONCE per remote node, to its "mirror" (n', l);
                                2. every rank forwards what arrived over the NIC to its 7 NVLink neighbours.
    NUMA 2-D / 3-D:             the same idea one level down, for PCIe boxes where GPUs under different CPU sockets
                                talk through the inter-socket link: (numa, local) or (node, numa, local).
    ..._ll variants:            the forwarding step spins on LL packets (tutorial 2, step 4) instead of a signal.

The index arithmetic every such kernel carries:   node = rank // local_world;  local = rank % local_world;
mirror on node n' = n' * local_world + local;  stage-2 peers = {node * local_world + l' for l' != local}.

WHAT IS LEFT ON 8 x B200. NVSwitch connects every GPU to every other GPU at the full NVLink-5 rate in ONE hop, and the
switch can replicate a store to all GPUs (multicast). There is no slow level to protect: a second hop would only add
latency and double the traffic of the forwarding GPU. So in this framework

  * the hierarchical MODE NAMES exist (code written against the reference keeps running) and all resolve to the flat
    one-hop kernels of tutorial 2: ``push3d`` / ``push_numa_2d`` -> P2P push, ``push_numa_2d_ll[_multinode]`` -> LL packets;
  * contexts refuse ``num_nodes != 1`` and the inter-node copy-engine producer raises NotImplementedError with the
    reason, instead of silently running something that is not what the name says;
  * the "one send per node, replicate at the destination" idea survives in hardware form: ``push_multimem`` sends the
    shard once and the SWITCH replicates it (tutorial 2, step 3).

This script runs every mode against the NCCL result, prints which kernel served it and its latency, and shows the
refusal of the inter-node entry points.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia import cp_engine_producer_all_gather_inter_node, create_fast_allgather_context, fast_allgather
from triton_dist.kernels.nvidia.low_latency_allgather import FastAllGatherContext
from triton_dist.utils import (dist_print, finalize_distributed, initialize_distributed, nvshmem_barrier_all_on_stream,
                               perf_func_max_over_ranks, symm_tensor)

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
gpu = torch.cuda.is_available()
dev = "cuda" if gpu else "cpu"

# ---- the 2-D arithmetic, for a pretend cluster of `nodes` x `local_world` (pure Python: nothing is sent)
local_world = max(world // 2, 1)
node, local = rank // local_world, rank % local_world
mirrors = [n * local_world + local for n in range(world // local_world) if n != node]
neighbours = [node * local_world + l for l in range(local_world) if l != local]
dist_print(f"rank {rank} seen as (node {node}, local {local}) of {world // local_world} x {local_world}: stage 1 -> mirrors "
           f"{mirrors}, stage 2 -> neighbours {neighbours}", allowed_ranks="all")

# ---- every mode name of the reference, served by the flat kernels
served_by = {"pull": "P2P loads", "push2d": "P2P push", "push3d": "P2P push", "push_numa_2d": "P2P push",
             "push2d_ll": "LL packets", "push_numa_2d_ll": "LL packets", "push_numa_2d_ll_multinode": "LL packets",
             "push2d_ll_multimem": "LL packets via multicast", "push_multimem": "multicast push"}
ctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=1 << 22)
n_words = 4096                                           # 16 KB per rank: small enough for the LL modes
buf = symm_tensor((world, n_words), torch.int32)
ref = torch.empty(world * n_words, dtype=torch.int32, device=dev)
mine = torch.full((n_words,), rank + 1, dtype=torch.int32, device=dev)
if world > 1:
    dist.all_gather_into_tensor(ref, mine)
else:
    ref.copy_(mine)
if gpu:
    nvshmem_barrier_all_on_stream()                      # all zero fills are done before the first push
for mode in (served_by if gpu else ["pull"]):
    buf[rank].copy_(mine)
    fast_allgather(buf, ctx, mode=mode)
    ok = torch.equal(buf.view(-1), ref)
    line = f"mode {mode:28s} -> {served_by[mode]:26s} {'ok' if ok else 'WRONG'}"
    if gpu:
        nvshmem_barrier_all_on_stream()                  # peers may still be comparing: do not refill yet
        _, t = perf_func_max_over_ranks(lambda: fast_allgather(buf, ctx, mode=mode), 20, 5)
        line += f"   {t * 1e3:6.1f} us"
    dist_print(line)
    assert ok

# ---- the entry points that would need a second node say so
try:
    cp_engine_producer_all_gather_inter_node()
except NotImplementedError as e:
    dist_print(f"cp_engine_producer_all_gather_inter_node: {e}")
try:
    FastAllGatherContext(rank=rank, node=0, num_ranks=world, num_nodes=2)
except NotImplementedError as e:
    dist_print(f"FastAllGatherContext(num_nodes=2): {e}")
if gpu:
    torch.cuda.synchronize()
ctx.finalize()
finalize_distributed()
