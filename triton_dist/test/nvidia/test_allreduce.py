"""AllReduce correctness + bus-bandwidth sweep (mirrors reference test/nvidia/test_allreduce.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_allreduce [--method two_shot_multimem] [--check]"""
import torch
import torch.distributed as dist

from triton_dist.kernels.allreduce import AllReduceMethod, get_allreduce_methods, to_allreduce_method
from triton_dist.kernels.nvidia.allreduce import all_reduce, create_allreduce_ctx
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--method", default="all", choices=["all", "auto"] + get_allreduce_methods())
    p.add_argument("--min-bytes", type=int, default=4096)
    p.add_argument("--max-bytes", type=int, default=64 * 2**20)
    args = p.parse_args()
    group, rank, world = setup()
    dtype = DTYPES[args.dtype]
    ctx = create_allreduce_ctx(min(max(args.max_bytes, 2**20) * 2 * world + 2**20, 1 << 30), rank, world, world)
    methods = get_allreduce_methods() if args.method == "all" else [args.method]
    ok = True
    sizes = []
    n = args.min_bytes
    while n <= args.max_bytes:
        sizes.append(n)
        n *= 4
    for name in methods:
        method = AllReduceMethod.Unknown if name == "auto" else to_allreduce_method(name)
        for nbytes in sizes:
            numel = nbytes // torch.empty((), dtype=dtype).element_size()
            for it in range(args.verify_iters):
                x = rand((numel,), dtype)
                skew(args.stress)
                y = all_reduce(x, method=method, ctx=ctx)
                ref = x.float()
                if world > 1:
                    dist.all_reduce(ref)
                ok = all_ok(check_close(y, ref, dtype, f"{name} {nbytes}B iter {it}"))
                if not ok:
                    break
            if not ok:
                break
            if ok and not args.check:
                x = rand((numel,), dtype)
                out = torch.empty_like(x)
                t = bench(lambda: all_reduce(x, out, method=method, ctx=ctx), args, flush_l2=False)
                ref = x.clone()
                t_nccl = bench(lambda: dist.all_reduce(ref), args, flush_l2=False) if world > 1 else float("nan")
                algo = nbytes / t / 1e6
                dist_print(f"allreduce {name:22s} {nbytes:>10d} B: {t * 1e3:8.1f} us  algbw {algo:7.1f} GB/s  "
                           f"busbw {algo * 2 * (world - 1) / world:7.1f} GB/s   NCCL {t_nccl * 1e3:8.1f} us")
        if not ok:
            break
    ctx.finalize()
    finish(ok)
