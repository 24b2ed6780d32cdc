"""Tutorial 8 — GEMM + ReduceScatter in one kernel (reference tutorials/08-overlapping-gemm-reduce-scatter.py).
Run:  bash scripts/launch.sh tutorials/08-overlapping-gemm-reduce-scatter.py

Row-parallel linear: every rank multiplies by its K-shard, the partial [M, N] results must be summed and rank r keeps
rows r*M/W... The kernel's epilogue stores each finished tile straight into the OWNER rank's memory over NVLink (tcgen05
accumulators -> registers -> peer store), tiles for other ranks first, publishes a flag per 128-row block, and the same
grid then reduces the W contributions of its own rows. Compare with cuBLAS followed by NCCL reduce_scatter."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia import create_gemm_rs_context, gemm_rs
from triton_dist.kernels.nvidia.gemm_reduce_scatter import torch_gemm_rs
from triton_dist.kernels.nvidia.gemm_rs_threadblock_swizzle import gemm_rs_tile_order
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
dev = "cuda" if torch.cuda.is_available() else "cpu"
dtype = torch.bfloat16 if dev == "cuda" else torch.float32
M, N, K = 8192, 4096, 14336  # Llama-3-8B down projection
a = torch.randn(M, K // world, device=dev).to(dtype) * 0.05
b = torch.randn(N, K // world, device=dev).to(dtype) * 0.05
ctx = create_gemm_rs_context(M, N, rank, world, world, dtype)
out = gemm_rs(a, b, ctx)
ref = torch_gemm_rs(a, b)
dist_print(f"rank {rank}: out {tuple(out.shape)}, max |diff| vs cuBLAS+NCCL {(out.float() - ref.float()).abs().max().item():.3f}", allowed_ranks="all")
dist_print(f"first tiles rank 0 computes (destination shard, m block, n block): {gemm_rs_tile_order(M, N, world, 0)[:4]}")
if dev == "cuda":
    _, t = perf_func_max_over_ranks(lambda: gemm_rs(a, b, ctx), 10, 3, flush_l2=True)
    _, tr = perf_func_max_over_ranks(lambda: torch_gemm_rs(a, b), 10, 3, flush_l2=True)
    dist_print(f"fused {t:.4f} ms ({2.0 * M * N * K / world / t / 1e9:.0f} TFLOP/s per GPU), cuBLAS + NCCL {tr:.4f} ms, speed-up {tr / t:.2f}x")
ctx.finalize()
finalize_distributed()
