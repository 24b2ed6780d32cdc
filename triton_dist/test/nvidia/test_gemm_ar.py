"""GEMM+AllReduce correctness / perf (mirrors reference test/nvidia/test_gemm_ar.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_gemm_ar 128 5120 25600 [--check] [--low-latency]"""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.gemm_allreduce import (create_gemm_ar_context, gemm_allreduce_op,
                                                       low_latency_gemm_allreduce_op, torch_gemm_ar)
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("M", type=int)
    p.add_argument("N", type=int)
    p.add_argument("K", type=int)
    p.add_argument("--low-latency", action="store_true")
    args = p.parse_args()
    group, rank, world = setup()
    dtype = DTYPES[args.dtype]
    k_loc = args.K // world
    B = rand((args.N, k_loc), dtype, 0.1)
    ctx = create_gemm_ar_context(None, rank, world, world, args.M, args.N, dtype)
    op = low_latency_gemm_allreduce_op if args.low_latency else gemm_allreduce_op
    ok = True
    for it in range(args.verify_iters):
        A = rand((args.M, k_loc), dtype)
        skew(args.stress)
        C = op(ctx, A, B)
        ref = torch_gemm_ar(A.float(), B.float())
        ok = all_ok(check_close(C, ref, dtype, f"gemm_ar iter {it}"))
        if not ok:
            break
    if ok and not args.check:
        A = rand((args.M, k_loc), dtype)
        t_fused = bench(lambda: op(ctx, A, B), args)
        t_ref = bench(lambda: torch_gemm_ar(A, B), args)
        dist_print(f"gemm_ar M={args.M} N={args.N} K={args.K} W={world}: fused {t_fused * 1e3:.1f} us, "
                   f"torch(cuBLAS+NCCL) {t_ref * 1e3:.1f} us, speedup {t_ref / t_fused:.2f}x")
    ctx.finalize()
    finish(ok)
