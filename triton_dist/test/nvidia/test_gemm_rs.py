"""GEMM+RS correctness / perf (mirrors reference test/nvidia/test_gemm_rs.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_gemm_rs 8192 4096 14336 [--check]
M global rows, N output columns, K GLOBAL reduction dim (sharded over ranks)."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs, torch_gemm_rs
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print, measured_peaks, NVLINK_GBS_MEASURED

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("M", type=int)
    p.add_argument("N", type=int)
    p.add_argument("K", type=int)
    p.add_argument("--block-n", type=int, default=0)
    p.add_argument("--alternate", action="store_true", help="two sizes back to back on one context, rank skew, no host sync")
    p.add_argument("--autotune", action="store_true", help="sweep the launch knobs under the contextual autotuner, then check the winner")
    args = p.parse_args()
    group, rank, world = setup()
    dtype = DTYPES[args.dtype]
    assert args.M % world == 0 and args.K % world == 0
    k_loc = args.K // world
    B = rand((args.N, k_loc), dtype, 0.1)
    ctx = create_gemm_rs_context(args.M, args.N, rank, world, world, dtype, BLOCK_N=args.block_n)
    ok = True
    for it in range(args.verify_iters):
        A = rand((args.M, k_loc), dtype)
        skew(args.stress)
        C = gemm_rs(A, B, ctx)
        ref = torch_gemm_rs(A.float(), B.float())  # fp32 ground truth
        ok = all_ok(check_close(C, ref, dtype, f"gemm_rs iter {it}"))
        if not ok:
            break
    if ok and args.alternate and (args.M // 2) % world == 0:
        ins, outs = [], []
        for it in range(8):
            ins.append(rand((args.M if it % 2 == 0 else args.M // 2, k_loc), dtype))
        torch.cuda.synchronize()
        dist.barrier()
        for it, a_in in enumerate(ins):
            if it % 3 == rank % 3:
                skew(True)
            outs.append(gemm_rs(a_in, B, ctx))
        for it, (a_in, c_out) in enumerate(zip(ins, outs)):
            ok = all_ok(check_close(c_out, torch_gemm_rs(a_in.float(), B.float()), dtype, f"gemm_rs alternate {it}")) and ok
    if ok and args.autotune:
        from triton_dist.autotuner import _REGISTRY, contextual_autotune

        A = rand(A.shape, dtype)
        C = contextual_autotune(is_dist=True, warmup=2, iters=5)(lambda: gemm_rs(A, B, ctx, autotune=True))()
        ok = all_ok(check_close(C, torch_gemm_rs(A.float(), B.float()), dtype, "gemm_rs autotuned"))
        dist_print(f"gemm_rs autotune winner: {list(_REGISTRY['gemm_rs'].best.values())}")
    if ok and not args.check:
        A = rand((args.M, k_loc), dtype)
        t_fused = bench(lambda: gemm_rs(A, B, ctx), args)
        t_ref = bench(lambda: torch_gemm_rs(A, B), args)
        flops = 2.0 * args.M * args.N * k_loc
        link_bytes = (world - 1) * (args.M // world) * args.N * A.element_size()
        pk = measured_peaks()
        roof_ms = max(flops / (pk["bf16_tflops"] * 1e9), link_bytes / (NVLINK_GBS_MEASURED * 1e6))
        dist_print(f"gemm_rs M={args.M} N={args.N} K={args.K} W={world}: fused {t_fused:.4f} ms "
                   f"({flops / t_fused / 1e9:.1f} TFLOP/s per GPU), torch(cuBLAS+NCCL) {t_ref:.4f} ms, "
                   f"speedup {t_ref / t_fused:.2f}x, roofline {roof_ms:.4f} ms -> {roof_ms / t_fused:.2f} of measured")
    ctx.finalize()
    finish(ok)
