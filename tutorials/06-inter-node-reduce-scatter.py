"""Tutorial 6 — the two-level reduce-scatter (reference tutorials/06-inter-node-reduce-scatter.py) on a flat node.
Run:  bash scripts/launch.sh tutorials/06-inter-node-reduce-scatter.py

The reference reduces inside a node first, then exchanges per-node partials over InfiniBand (kernel K14 + ring reduce).
With one NVSwitch node there is no second level; what remains useful from that design is fusing the *producer* into the
reduce-scatter: `gemm_rs` computes the partial tiles and scatters + reduces them inside ONE kernel. This tutorial shows
that both compositions give the same answer and times them."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia import create_gemm_rs_context, gemm_persistent, gemm_rs
from triton_dist.kernels.nvidia.reduce_scatter import create_reduce_scater_2d_ctx, reduce_scatter_2d_op
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
dev = "cuda" if torch.cuda.is_available() else "cpu"
dtype = torch.bfloat16 if dev == "cuda" else torch.float32
M, N, K = 1024 * world, 2048, 512 * world
a = torch.randn(M, K // world, device=dev).to(dtype) * 0.1
b = torch.randn(N, K // world, device=dev).to(dtype) * 0.1
rs_ctx = create_gemm_rs_context(M, N, rank, world, world, dtype)
sc_ctx = create_reduce_scater_2d_ctx(M, N, rank, world, world, dtype)
fused = gemm_rs(a, b, rs_ctx)
two_step = reduce_scatter_2d_op(gemm_persistent(a, b), sc_ctx)
err = (fused.float() - two_step.float()).abs().max().item()
dist_print(f"fused GEMM+RS vs GEMM then RS: max |diff| {err:.4f}")
if dev == "cuda":
    _, t1 = perf_func_max_over_ranks(lambda: gemm_rs(a, b, rs_ctx), 10, 3)
    _, t2 = perf_func_max_over_ranks(lambda: reduce_scatter_2d_op(gemm_persistent(a, b), sc_ctx), 10, 3)
    dist_print(f"one kernel {t1 * 1e3:.1f} us, two kernels {t2 * 1e3:.1f} us")
rs_ctx.finalize(); sc_ctx.finalize()
finalize_distributed()
