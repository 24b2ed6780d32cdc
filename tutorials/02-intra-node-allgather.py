"""Tutorial 2 — all-gather inside one NVSwitch node, three ways (reference tutorials/02-intra-node-allgather.py).
Run:  bash scripts/launch.sh tutorials/02-intra-node-allgather.py

1. copy engines + stream memory-ops: per-source `cuStreamWriteValue32` signals, the producer of the reference's AG+GEMM;
2. a push kernel over NVLS multicast: one `multimem.st` reaches all 8 GPUs (large messages);
3. the LL-packet kernel: 8-byte {data, flag} packets, no separate signal, lowest latency (small messages)."""
import torch
import torch.distributed as dist

from triton_dist.kernels.common_ops import CMP_GE, wait_eq
from triton_dist.kernels.nvidia import cp_engine_producer_all_gather_intra_node, create_fast_allgather_context, fast_allgather
from triton_dist.utils import (dist_print, finalize_distributed, initialize_distributed, nvshmem_barrier_all_on_stream,
                               nvshmem_create_tensors, symm_tensor)

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
if not torch.cuda.is_available():
    dist_print("this tutorial needs GPUs (peer mappings, copy engines)")
    finalize_distributed()
    raise SystemExit(0)
m, K = 1024, 4096
x = torch.full((m, K), float(rank), device="cuda", dtype=torch.bfloat16)

# (1) copy engines: every rank owns a symmetric [world*m, K] buffer; bufs[p] is peer p's instance mapped into my address space
bufs = nvshmem_create_tensors((world * m, K), torch.bfloat16, rank, world)
sigs = nvshmem_create_tensors((world,), torch.int32, rank, world)
bufs[rank][rank * m:(rank + 1) * m].copy_(x)
nvshmem_barrier_all_on_stream()
ag_stream = torch.cuda.Stream()
ag_stream.wait_stream(torch.cuda.current_stream())
cp_engine_producer_all_gather_intra_node(rank, world, x, bufs, ag_stream, sigs, signal_target=1)
for src in range(world):   # a consumer (e.g. a GEMM over shard `src`) waits on the stream, not on the host
    if src != rank:
        wait_eq(sigs[rank][src:src + 1], 1, cmp=CMP_GE)
ok1 = all(bool((bufs[rank][s * m:(s + 1) * m] == s).all()) for s in range(world))

# (2) + (3) kernels
ctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=world * m * K * 2)
big = symm_tensor((world, m * K), torch.bfloat16)
big[rank].fill_(rank)
fast_allgather(big, ctx, mode="push_multimem")
small = symm_tensor((world, 2048), torch.int32)
small[rank].fill_(rank)
fast_allgather(small, ctx, mode="push2d_ll")
torch.cuda.synchronize()
ok2 = all(bool((big[s] == s).all()) and bool((small[s] == s).all()) for s in range(world))
dist_print(f"rank {rank}: copy-engine all-gather {'ok' if ok1 else 'WRONG'}, kernel all-gathers {'ok' if ok2 else 'WRONG'}", allowed_ranks="all")
ctx.finalize()
finalize_distributed()
