"""TP_MoE layer: AG + grouped GEMM -> SiLU*up -> grouped GEMM + top-k reduce + RS against the per-expert loop oracle
(mirrors reference test/nvidia/test_tp_moe.py, test_ag_moe.py, test_moe_reduce_rs.py)."""
import torch
import torch.distributed as dist

from triton_dist.layers.nvidia import TP_MoE
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=1024, help="global tokens")
    p.add_argument("--hidden", type=int, default=1024)
    p.add_argument("--inter", type=int, default=512)
    p.add_argument("-E", type=int, default=16)
    p.add_argument("--topk", type=int, default=4)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    moe = TP_MoE(rank, world, group)
    gen = torch.Generator(device=dev).manual_seed(100 + rank)
    moe._init_random(args.hidden, args.inter, args.E, args.topk, dtype, gen, std=0.05)
    moe._init_ctx(max_M=args.M)
    moe._init_AR_ctx(args.M, None, dtype)
    m_loc = args.M // world
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(50 + it)
        x_full = rand((args.M, args.hidden), dtype)  # same on every rank (same seed)
        x_loc = x_full[rank * m_loc:(rank + 1) * m_loc].contiguous()
        skew(args.stress)
        out = moe.dist_triton_fwd(x_loc)
        ref_full = moe.torch_fwd(x_full)
        ref = ref_full[rank * m_loc:(rank + 1) * m_loc]
        good = check_close(out, ref.float(), dtype if dev == "cuda" else torch.float16, f"tp_moe(ag_rs) iter {it}")
        out_ar = moe.dist_triton_AR_fwd(x_full)
        good = good and check_close(out_ar, ref_full.float(), dtype if dev == "cuda" else torch.float16, f"tp_moe(AR) iter {it}")
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        x_loc = rand((m_loc, args.hidden), dtype)
        x_full = rand((args.M, args.hidden), dtype)
        t = bench(lambda: moe.dist_triton_fwd(x_loc), args)
        t_ref = bench(lambda: moe.torch_fwd(x_full), args)
        dist_print(f"tp_moe M={args.M} H={args.hidden} I={args.inter} E={args.E} topk={args.topk} W={world}: "
                   f"fused {t:.3f} ms, torch loop {t_ref:.3f} ms")
    moe.finalize()
    finish(ok)
