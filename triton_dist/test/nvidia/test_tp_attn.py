"""TP_Attn prefill + decode in every communication mode against the torch forward (mirrors reference
test/nvidia/test_tp_attn.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_tp_attn --bsz 8 --seq 128"""
import torch

from triton_dist.kernels.allreduce import AllReduceMethod
from triton_dist.layers.nvidia import TP_Attn
from triton_dist.layers.nvidia.tp_attn import _set_cos_sin_cache
from triton_dist.models import KV_Cache
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--bsz", type=int, default=0, help="global batch (default: 2 per rank)")
    p.add_argument("--seq", type=int, default=32)
    p.add_argument("--hidden", type=int, default=1024)
    p.add_argument("--heads", type=int, default=0)
    p.add_argument("--kv-heads", type=int, default=0)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    B = args.bsz or 2 * world
    heads = args.heads or max(8, world)
    kv_heads = args.kv_heads or max(2, world)
    D = 128
    attn = TP_Attn(rank, world, group, head_dim=D)
    gen = torch.Generator(device=dev).manual_seed(11)
    attn._init_random(args.hidden, heads, kv_heads, dtype, qk_norm=True, gen=gen, std=0.03)
    max_m = B * args.seq
    attn._init_ctx(max_M=max_m)
    attn._init_AR_ctx(max_m, AllReduceMethod.Unknown, dtype)
    attn._init_gemm_ar_ctx(max_m, dtype)
    inv_freq = 1.0 / (1e6 ** (torch.arange(0, D, 2, device=dev).float() / D))
    cos_sin = _set_cos_sin_cache(inv_freq, args.seq + 8)
    tol_dtype = dtype if dev == "cuda" else torch.float16

    def fresh_cache():
        return KV_Cache(num_layers=1, batch_size=B, max_length=args.seq + 8, kv_heads=kv_heads, head_dim=D, dtype=dtype, world_size=world)

    ok = True
    b_loc = B // world
    for it in range(args.verify_iters):
        torch.manual_seed(30 + it)
        x = rand((B, args.seq, args.hidden), dtype)
        pos = torch.arange(args.seq, device=dev).unsqueeze(0).expand(B, -1).contiguous()
        ref = attn.torch_fwd(x, pos, cos_sin, fresh_cache(), 0).float()
        skew(args.stress)
        good = check_close(attn.dist_triton_AR_fwd(x, pos, cos_sin, fresh_cache(), 0), ref, tol_dtype, f"tp_attn prefill AR iter {it}")
        good = check_close(attn.dist_triton_gemm_ar_fwd(x, pos, cos_sin, fresh_cache(), 0), ref, tol_dtype, f"tp_attn prefill gemm_ar iter {it}") and good
        x_loc = x[rank * b_loc:(rank + 1) * b_loc].contiguous()
        good = check_close(attn.dist_triton_fwd(x_loc, pos, cos_sin, fresh_cache(), 0), ref[rank * b_loc:(rank + 1) * b_loc], tol_dtype,
                           f"tp_attn prefill ag_rs iter {it}") and good
        # one decode step on top of a prefilled cache
        cache_ref, cache_new = fresh_cache(), fresh_cache()
        attn.torch_fwd(x, pos, cos_sin, cache_ref, 0)
        attn.dist_triton_AR_fwd(x, pos, cos_sin, cache_new, 0)
        xd = rand((B, 1, args.hidden), dtype)
        pd = torch.full((B, 1), args.seq, device=dev, dtype=torch.long)
        refd = attn.torch_fwd(xd, pd, cos_sin, cache_ref, 0).float()
        good = check_close(attn.dist_triton_AR_fwd(xd, pd, cos_sin, cache_new, 0), refd, tol_dtype, f"tp_attn decode AR iter {it}") and good
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        x = rand((B, args.seq, args.hidden), dtype)
        pos = torch.arange(args.seq, device=dev).unsqueeze(0).expand(B, -1).contiguous()
        cache = fresh_cache()
        t_ref = bench(lambda: attn.torch_fwd(x, pos, cos_sin, cache, 0), args)
        t_ar = bench(lambda: attn.dist_triton_AR_fwd(x, pos, cos_sin, cache, 0), args)
        t_gar = bench(lambda: attn.dist_triton_gemm_ar_fwd(x, pos, cos_sin, cache, 0), args)
        dist_print(f"tp_attn prefill B={B} S={args.seq} H={args.hidden} W={world}: torch {t_ref:.4f} ms | AR {t_ar:.4f} ms ({t_ref / t_ar:.2f}x) | "
                   f"gemm_ar {t_gar:.4f} ms ({t_ref / t_gar:.2f}x)")
    attn.finalize()
    finish(ok)
