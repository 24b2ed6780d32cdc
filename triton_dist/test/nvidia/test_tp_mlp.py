"""TP_MLP in every communication mode against the torch (NCCL) forward (mirrors reference test/nvidia/test_tp_mlp.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_tp_mlp --M 4096 --hidden 5120 --inter 25600"""
import torch

from triton_dist.kernels.allreduce import AllReduceMethod
from triton_dist.layers.nvidia import TP_MLP
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--M", type=int, default=1024, help="global tokens")
    p.add_argument("--hidden", type=int, default=1024)
    p.add_argument("--inter", type=int, default=2048)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    mlp = TP_MLP(rank, world, group)
    gen = torch.Generator(device=dev).manual_seed(7)  # same full weights on every rank; the layer keeps its shard
    mlp._init_random(args.hidden, args.inter, dtype, gen, std=0.03)
    mlp._init_ctx(max_M=args.M)
    mlp._init_AR_ctx(args.M, AllReduceMethod.Unknown, dtype)
    mlp._init_gemm_ar_ctx(args.M, dtype)
    m_loc = args.M // world
    ok = True
    tol_dtype = dtype if dev == "cuda" else torch.float16
    for it in range(args.verify_iters):
        torch.manual_seed(90 + it)
        x_full = rand((args.M, args.hidden), dtype)
        x_loc = x_full[rank * m_loc:(rank + 1) * m_loc].contiguous()
        ref_full = mlp.torch_fwd(x_full).float()
        skew(args.stress)
        good = check_close(mlp.dist_triton_fwd(x_loc), ref_full[rank * m_loc:(rank + 1) * m_loc], tol_dtype, f"tp_mlp ag_rs iter {it}")
        good = check_close(mlp.dist_triton_AR_fwd(x_full), ref_full, tol_dtype, f"tp_mlp AR iter {it}") and good
        good = check_close(mlp.dist_triton_gemm_ar_fwd(x_full), ref_full, tol_dtype, f"tp_mlp gemm_ar iter {it}") and good
        # the two halves on their own (reference test_tp_mlp.py checks ag_gemm and gemm_rs separately as well)
        good = check_close(mlp.dist_triton_ag_gemm(x_loc), mlp.torch_ag_gemm(x_loc).float(), tol_dtype, f"tp_mlp ag_gemm iter {it}") and good
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        x_loc = rand((m_loc, args.hidden), dtype)
        x_full = rand((args.M, args.hidden), dtype)
        t_ref = bench(lambda: mlp.torch_fwd(x_full), args)
        t_agrs = bench(lambda: mlp.dist_triton_fwd(x_loc), args)
        t_ar = bench(lambda: mlp.dist_triton_AR_fwd(x_full), args)
        t_gar = bench(lambda: mlp.dist_triton_gemm_ar_fwd(x_full), args)
        dist_print(f"tp_mlp M={args.M} H={args.hidden} I={args.inter} W={world}: torch {t_ref:.4f} ms | ag_gemm+gemm_rs {t_agrs:.4f} ms "
                   f"({t_ref / t_agrs:.2f}x) | gemm+AR {t_ar:.4f} ms ({t_ref / t_ar:.2f}x) | fused gemm_ar {t_gar:.4f} ms ({t_ref / t_gar:.2f}x)")
    mlp.finalize()
    finish(ok)
