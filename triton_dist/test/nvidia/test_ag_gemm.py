"""AG+GEMM correctness / perf (mirrors reference test/nvidia/test_ag_gemm.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_ag_gemm 8192 14336 4096 [--check]
M is the GLOBAL row count, N the GLOBAL column count (sharded over ranks), K the reduction dim."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.allgather_gemm import ag_gemm, create_ag_gemm_context, torch_ag_gemm
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print, measured_peaks, NVLINK_GBS_MEASURED

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("M", type=int)
    p.add_argument("N", type=int)
    p.add_argument("K", type=int)
    p.add_argument("--block-n", type=int, default=0)
    p.add_argument("--comm-sms", type=int, default=-1)
    p.add_argument("--alternate", action="store_true", help="two sizes back to back on one context, rank skew, no host sync")
    p.add_argument("--autotune", action="store_true", help="sweep the launch knobs under the contextual autotuner, then check the winner")
    args = p.parse_args()
    group, rank, world = setup()
    dtype = DTYPES[args.dtype]
    assert args.M % world == 0 and args.N % world == 0
    m_loc, n_loc = args.M // world, args.N // world
    B = rand((n_loc, args.K), dtype, 0.1)
    A0 = rand((m_loc, args.K), dtype)
    ctx = create_ag_gemm_context(A0, B, rank, world, max_M=args.M, num_local_ranks=world, BLOCK_N=args.block_n,
                                 num_comm_sms=args.comm_sms)
    ok = True
    for it in range(args.verify_iters):
        A = rand((m_loc, args.K), dtype)  # fresh inputs every iteration (stale-workspace detection)
        skew(args.stress)
        C = ag_gemm(A, B, ctx)
        ref = torch_ag_gemm(A.float(), B.float())
        ok = all_ok(check_close(C, ref, dtype, f"ag_gemm iter {it}"))
        if not ok:
            break
    if ok and args.alternate and m_loc % 2 == 0:
        # sizes alternate between calls: the other parity half of the workspace must still be intact when a fast rank
        # is already one call ahead (the halves are a fixed distance apart, not a size-dependent one)
        ins, outs = [], []
        for it in range(8):
            rows = m_loc if it % 2 == 0 else m_loc // 2
            ins.append(rand((rows, args.K), dtype))
        torch.cuda.synchronize()
        dist.barrier()
        for it, a_in in enumerate(ins):
            if it % 3 == rank % 3:
                skew(True)
            outs.append(ag_gemm(a_in, B, ctx))
        for it, (a_in, c_out) in enumerate(zip(ins, outs)):
            ok = all_ok(check_close(c_out, torch_ag_gemm(a_in.float(), B.float()), dtype, f"ag_gemm alternate {it}")) and ok
    if ok and args.autotune:
        from triton_dist.autotuner import _REGISTRY, contextual_autotune

        A = rand(A.shape, dtype)
        C = contextual_autotune(is_dist=True, warmup=2, iters=5)(lambda: ag_gemm(A, B, ctx, autotune=True))()
        ok = all_ok(check_close(C, torch_ag_gemm(A.float(), B.float()), dtype, "ag_gemm autotuned"))
        dist_print(f"ag_gemm autotune winner: {list(_REGISTRY['ag_gemm'].best.values())}")
    if ok and not args.check:
        A = rand((m_loc, args.K), dtype)
        t_fused = bench(lambda: ag_gemm(A, B, ctx), args)
        t_ref = bench(lambda: torch_ag_gemm(A, B), args)
        flops = 2.0 * args.M * n_loc * args.K
        link_bytes = (world - 1) * m_loc * args.K * A.element_size()
        pk = measured_peaks()
        roof_ms = max(flops / (pk["bf16_tflops"] * 1e9), link_bytes / (NVLINK_GBS_MEASURED * 1e6))
        dist_print(f"ag_gemm M={args.M} N={args.N} K={args.K} W={world}: fused {t_fused:.4f} ms "
                   f"({flops / t_fused / 1e9:.1f} TFLOP/s per GPU), torch(NCCL+cuBLAS) {t_ref:.4f} ms, "
                   f"speedup {t_ref / t_fused:.2f}x, roofline {roof_ms:.4f} ms -> {roof_ms / t_fused:.2f} of measured")
    ctx.finalize()
    finish(ok)
