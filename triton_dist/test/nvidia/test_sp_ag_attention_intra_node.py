"""SP all-gather attention vs an fp32 oracle on the un-sharded sequences (mirrors reference
test/nvidia/test_sp_ag_attention_intra_node.py)."""
import math

import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.sp_ag_attention_intra_node import (_global_pos, create_sp_ag_attention_context_intra_node,
                                                                   fused_sp_ag_attn_intra_node)
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--batch", type=int, default=2)
    p.add_argument("--seqlen", type=int, default=1024, help="GLOBAL sequence length per batch element")
    p.add_argument("--q-heads", type=int, default=8)
    p.add_argument("--kv-heads", type=int, default=2)
    p.add_argument("--no-zigzag", action="store_true")
    p.add_argument("--no-causal", action="store_true")
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    D, B, L = 128, args.batch, args.seqlen
    zig, causal = not args.no_zigzag, not args.no_causal
    assert L % (2 * world) == 0
    Ll = L // world
    ctx = create_sp_ag_attention_context_intra_node(B, args.q_heads, args.kv_heads, L, Ll, D, dtype, dtype, rank, world)
    cu_q = torch.arange(B + 1, dtype=torch.int32, device=dev) * Ll
    cu_k = torch.arange(B + 1, dtype=torch.int32, device=dev) * L
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(100 + it)  # identical full tensors on every rank; each takes its shard
        q_full = torch.randn(B, L, args.q_heads, D, device=dev).to(dtype)
        k_full = torch.randn(B, L, args.kv_heads, D, device=dev).to(dtype)
        v_full = torch.randn(B, L, args.kv_heads, D, device=dev).to(dtype)
        pos = _global_pos(torch.arange(Ll, device=dev), L, rank, world, zig)
        q = q_full[:, pos].reshape(B * Ll, args.q_heads, D).contiguous()
        k = k_full[:, pos].reshape(B * Ll, args.kv_heads, D).contiguous()
        v = v_full[:, pos].reshape(B * Ll, args.kv_heads, D).contiguous()
        skew(args.stress)
        out = fused_sp_ag_attn_intra_node(ctx, q, k, v, None, cu_q, cu_k, Ll, L, rank, world, causal, zig)
        g = args.q_heads // args.kv_heads
        qq = q_full.float().transpose(1, 2)
        kk = k_full.float().repeat_interleave(g, dim=2).transpose(1, 2)
        vv = v_full.float().repeat_interleave(g, dim=2).transpose(1, 2)
        ref_full = torch.nn.functional.scaled_dot_product_attention(qq, kk, vv, is_causal=causal).transpose(1, 2)
        ref = ref_full[:, pos].reshape(B * Ll, args.q_heads, D)
        ok = all_ok(check_close(out, ref, dtype if dev == "cuda" else torch.float16, f"sp_ag_attn iter {it}"))
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        t = bench(lambda: fused_sp_ag_attn_intra_node(ctx, q, k, v, None, cu_q, cu_k, Ll, L, rank, world, causal, zig), args)
        flops = 4.0 * B * args.q_heads * Ll * L * D * (0.5 if causal else 1.0)
        dist_print(f"sp_ag_attn B={B} L={L} Hq={args.q_heads} Hkv={args.kv_heads} W={world} zigzag={zig}: {t:.3f} ms "
                   f"({flops / t / 1e9:.1f} TFLOP/s per GPU)")
    ctx.finalize()
    finish(ok)
