"""Tutorial 7 — a tensor-parallel MLP with AG+GEMM and GEMM+RS (reference tutorials/07-08: overlapping AllGather GEMM /
GEMM ReduceScatter).  Run:  bash scripts/launch.sh tutorials/07-overlapping-allgather-gemm.py

Activations stay sharded on tokens between the two projections (Megatron sequence parallelism): the up-projection
gathers them INSIDE the GEMM kernel, the down-projection scatters and reduces its output INSIDE the GEMM kernel."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia import ag_gemm, create_ag_gemm_context, create_gemm_rs_context, gemm_rs
from triton_dist.kernels.nvidia.elementwise import silu_mul
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
dev = "cuda" if torch.cuda.is_available() else "cpu"
dtype = torch.bfloat16 if dev == "cuda" else torch.float32
M, H, I = 1024 * world, 1024, 4096
x = torch.randn(M // world, H, device=dev, dtype=dtype)                       # my token shard
w_gate_up = torch.randn(2 * I // world, H, device=dev, dtype=dtype) * 0.02    # column parallel (gate | up)
w_down = torch.randn(H, I // world, device=dev, dtype=dtype) * 0.02           # row parallel
ag_ctx = create_ag_gemm_context(x, w_gate_up, rank, world, max_M=M, num_local_ranks=world)
rs_ctx = create_gemm_rs_context(M, H, rank, world, world, dtype)
h = silu_mul(ag_gemm(x, w_gate_up, ag_ctx))        # [M, I/W]
y = gemm_rs(h, w_down, rs_ctx)                     # [M/W, H]
dist_print(f"rank {rank}: y {tuple(y.shape)} mean |y| = {y.float().abs().mean().item():.4f}", allowed_ranks="all")
ag_ctx.finalize(); rs_ctx.finalize()
finalize_distributed()
