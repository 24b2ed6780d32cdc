"""Reduce-scatter correctness / perf (mirrors reference test/nvidia/test_reduce_scatter.py)."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.reduce_scatter import create_reduce_scater_2d_ctx, reduce_scatter_2d_op, torch_reduce_scatter
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("M", type=int)
    p.add_argument("N", type=int)
    args = p.parse_args()
    group, rank, world = setup()
    dtype = DTYPES[args.dtype]
    ctx = create_reduce_scater_2d_ctx(args.M, args.N, rank, world, world, dtype)
    ok = True
    for it in range(args.verify_iters):
        x = rand((args.M, args.N), dtype)
        skew(args.stress)
        y = reduce_scatter_2d_op(x, ctx)
        ref = torch_reduce_scatter(x.float())
        ok = all_ok(check_close(y, ref, dtype, f"reduce_scatter iter {it}"))
        if not ok:
            break
    if ok and not args.check:
        x = rand((args.M, args.N), dtype)
        t = bench(lambda: reduce_scatter_2d_op(x, ctx), args, flush_l2=False)
        t_ref = bench(lambda: torch_reduce_scatter(x), args, flush_l2=False)
        nbytes = x.numel() * x.element_size()
        dist_print(f"reduce_scatter [{args.M},{args.N}] W={world}: {t * 1e3:.1f} us busbw "
                   f"{nbytes * (world - 1) / world / t / 1e6:.1f} GB/s, NCCL {t_ref * 1e3:.1f} us")
    ctx.finalize()
    finish(ok)
