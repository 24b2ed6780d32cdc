"""Distributed flash-decode: KV sharded over ranks on the sequence dim (mirrors reference test/nvidia/test_sp_decode_attn.py)."""
import math

import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.flash_decode import torch_paged_decode
from triton_dist.layers.nvidia import SpGQAFlashDecodeAttention
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--batch", type=int, default=1)
    p.add_argument("--q-heads", type=int, default=32)
    p.add_argument("--kv-heads", type=int, default=8)
    p.add_argument("--kv-len", type=int, default=8192, help="GLOBAL kv length")
    p.add_argument("--page", type=int, default=16)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    D, B = 128, args.batch
    loc = args.kv_len // world
    pages = (loc + args.page - 1) // args.page
    scale = 1 / math.sqrt(D)
    layer = SpGQAFlashDecodeAttention(rank, 0, world, 1, args.q_heads, args.kv_heads, D, D, page_size=args.page, scale=scale,
                                      max_allowed_batch=B)
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(11 + it)  # same q on all ranks
        q = torch.randn(B, args.q_heads, D, device=dev).to(dtype)
        torch.manual_seed(1000 * rank + it)
        kc = torch.randn(B * pages, args.page, args.kv_heads, D, device=dev).to(dtype)
        vc = torch.randn(B * pages, args.page, args.kv_heads, D, device=dev).to(dtype)
        bt = torch.arange(B * pages, device=dev, dtype=torch.int32).view(B, pages)
        lens_loc = torch.full((B,), loc - rank, dtype=torch.int32, device=dev)  # ragged shards
        all_lens = [torch.zeros_like(lens_loc) for _ in range(world)]
        if world > 1:
            dist.all_gather(all_lens, lens_loc)
        else:
            all_lens = [lens_loc]
        gl = torch.stack(all_lens)
        skew(args.stress)
        out = layer(q, kc, vc, gl, bt)
        # oracle: gather every shard and attend over the concatenation
        ks, vs = [torch.zeros_like(kc) for _ in range(world)], [torch.zeros_like(vc) for _ in range(world)]
        if world > 1:
            dist.all_gather(ks, kc)
            dist.all_gather(vs, vc)
        else:
            ks, vs = [kc], [vc]
        parts = []
        for s in range(world):
            o, lse = torch_paged_decode(q, ks[s], vs[s], gl[s], bt, scale)
            parts.append((o, lse))
        lse_all = torch.stack([p_[1] for p_ in parts])
        w = torch.softmax(lse_all, dim=0)
        ref = sum(w[s][..., None] * parts[s][0] for s in range(world))
        ok = all_ok(check_close(out, ref, dtype if dev == "cuda" else torch.float16, f"sp decode iter {it}"))
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        t = bench(lambda: layer(q, kc, vc, gl, bt), args, flush_l2=False)
        kv_bytes = 2 * B * loc * args.kv_heads * D * 2
        dist_print(f"sp_flash_decode B={B} Hq={args.q_heads} Hkv={args.kv_heads} global_kv={args.kv_len} W={world}: "
                   f"{t * 1e3:.1f} us, local KV bandwidth {kv_bytes / t / 1e9:.2f} TB/s")
    layer.finalize()
    finish(ok)
