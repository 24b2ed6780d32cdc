"""AG+GEMM / GEMM+RS sweep over the shapes of the reference's headline charts (mirrors reference
python/triton_dist/benchmark/bench_allgather_gemm.py and README.md:191-201). Prints, per shape, the fused time, the
NCCL+cuBLAS time, the speed-up and the achieved fraction of the roofline (max of compute at the measured cuBLAS peak and
NVLink bytes at the measured per-direction bandwidth). Device-timed, L2 flushed, max over ranks.
    bash scripts/launch.sh -m triton_dist.benchmark.bench_allgather_gemm [--op ag_gemm|gemm_rs|both]"""
import argparse
import json

import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.allgather_gemm import ag_gemm, create_ag_gemm_context, torch_ag_gemm
from triton_dist.kernels.nvidia.comm_perf_model import fused_op_roofline_ms
from triton_dist.kernels.nvidia.gemm_perf_model import get_tensorcore_tflops
from triton_dist.kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs, torch_gemm_rs
from triton_dist.test.common import finish, rand, setup
from triton_dist.utils import dist_print, perf_func_max_over_ranks

# (M, N, K): N / K are the GLOBAL sharded dims (reference asset/ag-gemm-intra-node.png, gemm-rs-intranode-perf.png)
AG_SHAPES = [(8192, 11008, 4096), (8192, 14336, 4096), (8192, 28672, 8192), (8192, 53248, 16384), (8192, 29568, 8192)]
RS_SHAPES = [(8192, 4096, 11008), (8192, 4096, 14336), (8192, 3584, 14336), (8192, 4608, 36864), (8192, 8192, 28672), (8192, 8192, 29568)]

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="both", choices=["ag_gemm", "gemm_rs", "both"])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    group, rank, world = setup()
    dtype = torch.bfloat16
    peak = get_tensorcore_tflops(dtype)
    rows = []
    if args.op in ("ag_gemm", "both"):
        for (M, N, K) in AG_SHAPES:
            if N % (8 * world) or M % world:
                continue
            A, B = rand((M // world, K), dtype), rand((N // world, K), dtype, 0.05)
            ctx = create_ag_gemm_context(A, B, rank, world, max_M=M, num_local_ranks=world)
            _, t = perf_func_max_over_ranks(lambda: ag_gemm(A, B, ctx), args.iters, args.warmup, flush_l2=True)
            _, t0 = perf_func_max_over_ranks(lambda: torch_ag_gemm(A, B), args.iters, args.warmup, flush_l2=True)
            flops = 2.0 * M * (N // world) * K
            roof = fused_op_roofline_ms(flops, (world - 1) * (M // world) * K * 2, peak)
            rows.append(dict(op="ag_gemm", M=M, N=N, K=K, world=world, fused_ms=t, nccl_cublas_ms=t0, speedup=t0 / t,
                             tflops_per_gpu=flops / t / 1e9, roofline_frac=roof / t))
            ctx.finalize()
    if args.op in ("gemm_rs", "both"):
        for (M, N, K) in RS_SHAPES:
            if K % (8 * world) or M % world:
                continue
            A, B = rand((M, K // world), dtype), rand((N, K // world), dtype, 0.05)
            ctx = create_gemm_rs_context(M, N, rank, world, world, dtype)
            _, t = perf_func_max_over_ranks(lambda: gemm_rs(A, B, ctx), args.iters, args.warmup, flush_l2=True)
            _, t0 = perf_func_max_over_ranks(lambda: torch_gemm_rs(A, B), args.iters, args.warmup, flush_l2=True)
            flops = 2.0 * M * N * (K // world)
            roof = fused_op_roofline_ms(flops, (world - 1) * (M // world) * N * 2, peak)
            rows.append(dict(op="gemm_rs", M=M, N=N, K=K, world=world, fused_ms=t, nccl_cublas_ms=t0, speedup=t0 / t,
                             tflops_per_gpu=flops / t / 1e9, roofline_frac=roof / t))
            ctx.finalize()
    for r in rows:
        dist_print(f"{r['op']:8s} m{r['M']} n{r['N']} k{r['K']} W={r['world']}: fused {r['fused_ms']:.4f} ms, NCCL+cuBLAS {r['nccl_cublas_ms']:.4f} ms, "
                   f"speed-up {r['speedup']:.2f}x, {r['tflops_per_gpu']:.0f} TFLOP/s/GPU, {r['roofline_frac']:.2f} of roofline")
    if args.json and rank == 0:
        with open(args.json, "w") as f:
            json.dump(rows, f, indent=1)
    finish(True)
