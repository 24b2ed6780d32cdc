"""EP dispatch / combine round trip (mirrors reference test/nvidia/test_ep_a2a.py: dispatch -> combine == input * topk,
plus received rows checked against an all_to_all oracle).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_ep_a2a -M 4096 -N 7168 -G 256 --topk 8"""
import torch
import torch.distributed as dist

from triton_dist.layers.nvidia import EPAll2AllLayer
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=512, help="tokens per rank")
    p.add_argument("-N", type=int, default=1024, help="hidden")
    p.add_argument("-G", type=int, default=32, help="total experts")
    p.add_argument("--topk", type=int, default=4)
    p.add_argument("--fp8", action="store_true")
    args = p.parse_args()
    group, rank, world = setup()
    dtype = DTYPES[args.dtype] if torch.cuda.is_available() else torch.float32
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    layer = EPAll2AllLayer(group, args.M, args.N, args.topk, rank, args.G, world, world, dtype=dtype, fp8_dispatch=args.fp8)
    epr = args.G // world
    ok = True
    for it in range(args.verify_iters):
        x = rand((args.M, args.N), dtype)
        scores = torch.rand(args.M, args.G, device=dev)
        ids = scores.topk(args.topk, dim=-1).indices.int()
        skew(args.stress)
        buf, counts, n = layer.dispatch(x, ids, exact=False)
        n_recv = int(n.item())
        # oracle for the counts: how many pairs target my experts, from all ranks
        mine = ((ids // epr) == rank).sum().reshape(1).long()
        tot = [torch.zeros_like(mine) for _ in range(world)]
        if world > 1:
            dist.all_gather(tot, mine)
            # every rank r contributes its pairs destined to rank `rank`
            per_dst = torch.stack([((ids // epr) == d).sum() for d in range(world)]).long()
            allp = [torch.zeros_like(per_dst) for _ in range(world)]
            dist.all_gather(allp, per_dst)
            expect = int(sum(int(a[rank]) for a in allp))
        else:
            expect = int(mine.item())
        good = n_recv == expect and int(counts.sum()) == expect
        if not good:
            dist_print(f"iter {it}: received {n_recv}, expected {expect}", allowed_ranks="all", prefix=True)
        if args.fp8 and torch.cuda.is_available():
            rows = buf[:n_recv].view(torch.float8_e4m3fn).float().view(n_recv, args.N // 128, 128)
            deq = (rows * layer.dispatch_scales()[:n_recv, :, None]).view(n_recv, args.N).to(dtype)
        else:
            deq = buf[:n_recv].to(dtype)
        # identity "expert": combine must return x * topk (reference test_ep_a2a.py:370)
        out = layer.combine(deq)
        ref = x.float() * args.topk
        tol = (0.15 if args.fp8 else 2e-2) * float(ref.abs().max())
        err = float((out.float() - ref).abs().max())
        if err > tol:
            good = False
            dist_print(f"iter {it}: round trip max err {err:.4g} > {tol:.4g}", allowed_ranks="all", prefix=True)
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and torch.cuda.is_available():
        x = rand((args.M, args.N), dtype)
        ids = torch.rand(args.M, args.G, device=dev).topk(args.topk, dim=-1).indices.int()
        t_d = bench(lambda: layer.dispatch(x, ids, exact=False), args, flush_l2=False)
        eo = layer.ctx.expert_out
        t_c = bench(lambda: layer.combine(eo), args, flush_l2=False)
        nbytes = args.M * args.topk * args.N * (1 if args.fp8 else 2)
        dist_print(f"ep_a2a M={args.M} hidden={args.N} experts={args.G} topk={args.topk} W={world} fp8={args.fp8}: "
                   f"dispatch {t_d * 1e3:.1f} us ({nbytes / t_d / 1e6:.1f} GB/s per rank), combine {t_c * 1e3:.1f} us")
    layer.finalize()
    finish(ok)
