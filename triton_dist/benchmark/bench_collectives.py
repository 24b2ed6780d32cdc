"""Bus-bandwidth sweep of AllReduce / AllGather / ReduceScatter, 4 KB - 1 GB, vs NCCL (BASELINE.json config 5).
    bash scripts/launch.sh -m triton_dist.benchmark.bench_collectives [--max-bytes 1073741824]
busbw: allreduce 2(W-1)/W * bytes / t, allgather & reduce-scatter (W-1)/W * total bytes / t (NCCL-tests convention)."""
import argparse
import json

import torch
import torch.distributed as dist

from triton_dist.kernels.allreduce import AllReduceMethod
from triton_dist.kernels.nvidia.allreduce import all_reduce, create_allreduce_ctx
from triton_dist.kernels.nvidia.low_latency_allgather import create_fast_allgather_context, fast_allgather
from triton_dist.kernels.nvidia.reduce_scatter import create_reduce_scater_2d_ctx, reduce_scatter_2d_op
from triton_dist.test.common import finish, setup
from triton_dist.utils import dist_print, perf_func_max_over_ranks, symm_tensor

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", type=int, default=4096)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    group, rank, world = setup()
    actx = create_allreduce_ctx(1 << 30, rank, world, world)
    gctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=1 << 22)
    rows = []
    n = args.min_bytes
    while n <= args.max_bytes:
        graph = n <= (1 << 20)  # latency-bound sizes are timed inside a CUDA graph (no host launch overhead)
        x = torch.randn(n // 2, device="cuda").bfloat16()
        out = torch.empty_like(x)
        ref = x.clone()
        _, t = perf_func_max_over_ranks(lambda: all_reduce(x, out, AllReduceMethod.Unknown, actx), args.iters, 3, use_graph=graph)
        _, tn = perf_func_max_over_ranks(lambda: dist.all_reduce(ref), args.iters, 3, use_graph=graph)
        rows.append(dict(op="allreduce", bytes=n, us=t * 1e3, nccl_us=tn * 1e3, busbw=2 * (world - 1) / world * n / t / 1e6,
                         nccl_busbw=2 * (world - 1) / world * n / tn / 1e6))
        if n <= (1 << 28):
            buf = symm_tensor((world, n // world // 4), torch.int32)
            flat = torch.empty(n // 4, dtype=torch.int32, device="cuda")
            mode = "push2d_ll_multimem" if n <= (1 << 18) else "push_multimem"
            _, t = perf_func_max_over_ranks(lambda: fast_allgather(buf, gctx, mode=mode), args.iters, 3, use_graph=graph)
            _, tn = perf_func_max_over_ranks(lambda: dist.all_gather_into_tensor(flat, buf[rank]), args.iters, 3, use_graph=graph)
            rows.append(dict(op="allgather", bytes=n, us=t * 1e3, nccl_us=tn * 1e3, busbw=(world - 1) / world * n / t / 1e6,
                             nccl_busbw=(world - 1) / world * n / tn / 1e6))
            from triton_dist.runtime import get_heap

            get_heap().free(buf)
            M = max(world * 8, n // 2 // 4096 // world * world)
            if M * 4096 * 2 == n:
                rctx = create_reduce_scater_2d_ctx(M, 4096, rank, world, world, torch.bfloat16)
                xin = torch.randn(M, 4096, device="cuda").bfloat16()
                o2 = torch.empty(M // world, 4096, device="cuda", dtype=torch.bfloat16)
                _, t = perf_func_max_over_ranks(lambda: reduce_scatter_2d_op(xin, rctx, o2), args.iters, 3)
                _, tn = perf_func_max_over_ranks(lambda: dist.reduce_scatter_tensor(o2, xin), args.iters, 3)
                rows.append(dict(op="reduce_scatter", bytes=n, us=t * 1e3, nccl_us=tn * 1e3, busbw=(world - 1) / world * n / t / 1e6,
                                 nccl_busbw=(world - 1) / world * n / tn / 1e6))
                rctx.finalize()
        n *= 4
    for r in rows:
        dist_print(f"{r['op']:14s} {r['bytes']:>11d} B: {r['us']:9.1f} us  busbw {r['busbw']:7.1f} GB/s | NCCL {r['nccl_us']:9.1f} us  {r['nccl_busbw']:7.1f} GB/s")
    if args.json and rank == 0:
        with open(args.json, "w") as f:
            json.dump(rows, f, indent=1)
    actx.finalize()
    gctx.finalize()
    finish(True)
