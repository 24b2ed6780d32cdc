"""AllGather + grouped GEMM (MoE-TP first GEMM) against a per-expert torch loop (mirrors reference test_ag_moe.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_ag_moe -M 8192 --hidden 4096 --inter 2048 -E 32 --topk 4"""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.allgather_group_gemm import ag_group_gemm, create_ag_group_gemm_context
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=1024, help="global tokens")
    p.add_argument("--hidden", type=int, default=1024)
    p.add_argument("--inter", type=int, default=1024, help="FFN dim (sharded over ranks)")
    p.add_argument("-E", type=int, default=16)
    p.add_argument("--topk", type=int, default=4)
    p.add_argument("--overlap", action="store_true", help="stage-gated / column-chunked schedule (transfer overlaps the grouped GEMM)")
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    m_loc, n_loc = args.M // world, args.inter // world
    ctx = create_ag_group_gemm_context(args.M, args.hidden, args.E, args.topk, rank, world, dtype=dtype, overlap=args.overlap)
    torch.manual_seed(5 + rank)
    w = rand((args.E, n_loc, args.hidden), dtype) * 0.1
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(200 + it * 31 + rank)
        x = rand((m_loc, args.hidden), dtype)
        ids = torch.rand(m_loc, args.E, device=dev).topk(args.topk, dim=-1).indices.int()
        skew(args.stress)
        y, sort = ag_group_gemm(x, w, ids, ctx)
        xs = [torch.empty_like(x) for _ in range(world)]
        idl = [torch.empty_like(ids) for _ in range(world)]
        if world > 1:
            dist.all_gather(xs, x)
            dist.all_gather(idl, ids)
        else:
            xs, idl = [x], [ids]
        fx, fid = torch.cat(xs), torch.cat(idl).long()
        rows = sort.pair_to_row.long().view(args.M, args.topk)
        good = True
        for k in range(args.topk):
            ref = torch.einsum("th,tnh->tn", fx.float(), w[fid[:, k]].float())
            good = check_close(y[rows[:, k]], ref, dtype if dev == "cuda" else torch.float16, f"ag_group_gemm iter {it} k={k}") and good
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        x = rand((m_loc, args.hidden), dtype)
        ids = torch.rand(m_loc, args.E, device=dev).topk(args.topk, dim=-1).indices.int()
        t = bench(lambda: ag_group_gemm(x, w, ids, ctx), args)
        fl = 2.0 * args.M * args.topk * n_loc * args.hidden
        dist_print(f"ag_group_gemm M={args.M} H={args.hidden} I/W={n_loc} E={args.E} topk={args.topk} W={world}: {t:.4f} ms "
                   f"({fl / t / 1e9:.1f} TFLOP/s per GPU)")
    ctx.finalize()
    finish(ok)
