"""Tutorial 5 — reduce-scatter inside one node (reference tutorials/05-intra-node-reduce-scatter.py).
Run:  bash scripts/launch.sh tutorials/05-intra-node-reduce-scatter.py

Every rank holds a full [M, N] partial sum; rank r wants sum over ranks of rows [r*M/W, (r+1)*M/W). The kernel lets
rank r read its segment from all peers through NVLS `multimem.ld_reduce` (the switch adds the 8 copies in fp32 and
returns one), or with plain P2P loads when no multicast binding exists; the ring variants of the reference (PCIe boxes)
are aliases of the same kernel on NVSwitch."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.reduce_scatter import create_reduce_scater_2d_ctx, reduce_scatter_2d_op
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
dev = "cuda" if torch.cuda.is_available() else "cpu"
dtype = torch.bfloat16 if dev == "cuda" else torch.float32
M, N = 1024 * world, 4096
x = torch.randn(M, N, device=dev).to(dtype)
ctx = create_reduce_scater_2d_ctx(M, N, rank, world, world, dtype)
out = reduce_scatter_2d_op(x, ctx)
ref = x.float().clone()
if world > 1:
    dist.all_reduce(ref)
ref = ref[rank * (M // world):(rank + 1) * (M // world)]
err = (out.float() - ref).abs().max().item()
dist_print(f"rank {rank}: reduce-scatter [{M}, {N}] -> {tuple(out.shape)}, max |err| {err:.4f} (|ref| max {ref.abs().max().item():.2f})",
           allowed_ranks="all")
ctx.finalize()
finalize_distributed()
