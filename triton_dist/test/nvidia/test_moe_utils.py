"""MoE index preparation and top-k reduce against torch (mirrors reference test/nvidia/test_moe_utils.py):
moe_align_sort (native single-CTA sort) vs torch_align_sort, gather rows, grouped GEMM, weighted top-k reduce."""
import torch

from triton_dist.kernels.nvidia.moe_utils import (BLOCK_M, grouped_gemm, moe_align_sort, moe_gather_rows, reduce_topk,
                                                  torch_align_sort)
from triton_dist.test.common import DTYPES, all_ok, base_parser, check_close, finish, rand, setup
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-T", type=int, default=777, help="tokens")
    p.add_argument("-E", type=int, default=24)
    p.add_argument("--topk", type=int, default=6)
    p.add_argument("--hidden", type=int, default=512)
    p.add_argument("--inter", type=int, default=384)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(1000 + it + rank)
        T = args.T + 13 * it
        # skewed routing: some experts get nothing, some get a lot
        logits = torch.randn(T, args.E, device=dev) + torch.linspace(-2, 2, args.E, device=dev)
        ids = logits.topk(args.topk, dim=-1).indices
        wts = torch.softmax(torch.randn(T, args.topk, device=dev), -1)
        s = moe_align_sort(ids, args.E)
        ref = torch_align_sort(ids, args.E)
        good = bool(torch.equal(s.meta.cpu(), ref.meta.cpu()))
        rows = int(ref.meta[0])
        good = good and bool(torch.equal(s.block_expert[:rows // BLOCK_M].cpu(), ref.block_expert[:rows // BLOCK_M].cpu()))
        # the order inside an expert is free; check the layout instead: every pair sits in a block of its expert
        be = s.block_expert.long()[(s.pair_to_row.long() // BLOCK_M)]
        good = good and bool(torch.equal(be.cpu(), ids.reshape(-1).long().cpu()))
        good = good and bool(torch.equal(s.sorted_pair_ids[s.pair_to_row.long()].cpu(), torch.arange(T * args.topk, dtype=torch.int32)))
        if not good:
            dist_print(f"iter {it}: align/sort layout mismatch", allowed_ranks="all", prefix=True)
        x = rand((T, args.hidden), dtype)
        w = rand((args.E, args.inter, args.hidden), dtype) * 0.1
        a = moe_gather_rows(x, s)
        y = grouped_gemm(a, w, s)
        out = reduce_topk(y, s, wts)
        refo = torch.zeros(T, args.inter, device=dev)
        for k in range(args.topk):
            refo += wts[:, k:k + 1] * torch.einsum("th,tnh->tn", x.float(), w[ids[:, k]].float())
        good = check_close(out, refo, dtype if dev == "cuda" else torch.float16, f"moe gather+grouped_gemm+reduce_topk iter {it}") and good
        ok = all_ok(good)
        if not ok:
            break
    finish(ok)
