"""Low-latency all-to-all vs torch all_to_all_single (mirrors reference test/nvidia/test_all_to_all.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_all_to_all -M 128 -N 7168 -G 256 --topk 8 [--with-scale]"""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.low_latency_all_to_all import all_to_all_post_process, create_all_to_all_context, fast_all_to_all
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, finish, rand, setup, skew
from triton_dist.utils import all_to_all_compat, dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=128, help="tokens per rank (before top-k expansion)")
    p.add_argument("-N", type=int, default=1024)
    p.add_argument("-G", type=int, default=32)
    p.add_argument("--topk", type=int, default=4)
    p.add_argument("--with-scale", action="store_true")
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    epr = args.G // world
    max_m = args.M * args.topk
    ctx = create_all_to_all_context(max_m, args.N, rank, args.G, world, epr, dtype)

    def make_inputs():
        ids = torch.rand(args.M, args.G, device=dev).topk(args.topk, dim=-1).indices.reshape(-1)
        order = torch.argsort(ids, stable=True)
        x = rand((args.M, args.N), dtype)
        send = x[order // args.topk].contiguous()
        splits = torch.bincount(ids, minlength=args.G).int()
        cumsum = torch.cat([torch.zeros(1, dtype=torch.int32, device=dev), torch.cumsum(splits, 0).int()])
        scale = torch.rand(max_m, device=dev) if args.with_scale else None
        return send, splits, cumsum, scale

    ok = True
    for it in range(args.verify_iters):
        send, splits, cumsum, scale = make_inputs()
        skew(args.stress)
        rsplits, rdata, rscale = fast_all_to_all(ctx, send, cumsum, scale)
        out, out_scale = all_to_all_post_process(ctx, rsplits, rdata, rscale)
        # oracle
        in_counts = splits.view(world, epr).sum(1)
        all_counts = [torch.zeros_like(in_counts) for _ in range(world)]
        if world > 1:
            dist.all_gather(all_counts, in_counts)
        else:
            all_counts = [in_counts]
        out_counts = [int(c[rank]) for c in all_counts]
        ref = torch.empty((sum(out_counts), args.N), dtype=dtype, device=dev)
        if world > 1:
            all_to_all_compat(list(ref.split(out_counts)), list(send[:int(in_counts.sum())].split([int(c) for c in in_counts])))
        else:
            ref.copy_(send)
        good = out.shape == ref.shape and torch.equal(out, ref)
        if args.with_scale and good:
            sref = torch.empty((sum(out_counts),), dtype=torch.float32, device=dev)
            if world > 1:
                all_to_all_compat(list(sref.split(out_counts)), list(scale[:int(in_counts.sum())].split([int(c) for c in in_counts])))
            else:
                sref.copy_(scale[:sref.numel()])
            good = torch.equal(out_scale, sref)
        if not good:
            dist_print(f"iter {it}: mismatch", allowed_ranks="all", prefix=True)
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        send, splits, cumsum, scale = make_inputs()
        t = bench(lambda: fast_all_to_all(ctx, send, cumsum, scale), args, flush_l2=False)
        dist_print(f"fast_all_to_all {args.M} tok/rank topk {args.topk} hidden {args.N} W={world}: {t * 1e3:.1f} us")
    ctx.finalize()
    finish(ok)
