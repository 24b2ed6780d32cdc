"""Grouped GEMM + top-k reduce + reduce-scatter (MoE-TP second GEMM) against torch (mirrors reference
test_moe_reduce_rs.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_moe_reduce_rs -M 8192 --hidden 4096 --inter 2048"""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.moe_reduce_rs import create_moe_rs_context, run_moe_reduce_rs
from triton_dist.kernels.nvidia.moe_utils import moe_align_sort
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
    i_loc, m_loc = args.inter // world, args.M // world
    ctx = create_moe_rs_context(rank, world, world, args.M, args.hidden, args.E, args.topk, dtype, overlap=args.overlap)
    torch.manual_seed(9 + rank)
    w = rand((args.E, args.hidden, i_loc), dtype) * 0.1  # this rank's FFN shard of every expert's down projection
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(400 + it)  # routing and weights are global: same on every rank
        ids = torch.rand(args.M, args.E, device=dev).topk(args.topk, dim=-1).indices.int()
        wts = torch.softmax(torch.randn(args.M, args.topk, device=dev), -1)
        sort = moe_align_sort(ids, args.E)
        torch.manual_seed(500 + it * 7 + rank)
        h_pairs = rand((args.M * args.topk, i_loc), dtype)  # activations of every (token, k) pair on this FFN shard
        h_sorted = torch.zeros(sort.max_rows, i_loc, device=dev, dtype=dtype)
        h_sorted[sort.pair_to_row.long()] = h_pairs
        skew(args.stress)
        out = run_moe_reduce_rs(h_sorted, w, sort, wts, ctx)
        hp = h_pairs.view(args.M, args.topk, i_loc).float()
        full = torch.zeros(args.M, args.hidden, device=dev)
        for k in range(args.topk):
            full += wts[:, k:k + 1] * torch.einsum("ti,thi->th", hp[:, k], w[ids[:, k].long()].float())
        if world > 1:
            dist.all_reduce(full)
        good = check_close(out, full[rank * m_loc:(rank + 1) * m_loc], dtype if dev == "cuda" else torch.float16, f"moe_reduce_rs iter {it}")
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        t = bench(lambda: run_moe_reduce_rs(h_sorted, w, sort, wts, ctx), args)
        dist_print(f"moe_reduce_rs M={args.M} H={args.hidden} I/W={i_loc} E={args.E} topk={args.topk} W={world}: {t:.4f} ms")
    ctx.finalize()
    finish(ok)
