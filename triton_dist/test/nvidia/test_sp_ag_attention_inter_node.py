"""Counterpart of the reference's test_sp_ag_attention_inter_node.py: the inter-node entry points on a single NVSwitch node
(they forward to the intra-node kernel; asking for more than one node fails clearly)."""
import torch

from triton_dist.kernels.nvidia import create_sp_ag_attention_context_inter_node, fused_sp_ag_attn_inter_node
from triton_dist.kernels.nvidia.sp_ag_attention_intra_node import _global_pos
from triton_dist.test.common import DTYPES, all_ok, base_parser, check_close, finish, setup

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--seqlen", type=int, default=512)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    B, Hq, Hkv, D, L = 2, 8, 2, 128, args.seqlen * world  # same shape class as the intra-node test
    Ll = L // world
    ctx = create_sp_ag_attention_context_inter_node(B, Hq, Hkv, L, Ll, D, dtype, dtype, rank, rank, world, world)
    try:
        create_sp_ag_attention_context_inter_node(B, Hq, Hkv, L, Ll, D, dtype, dtype, rank, rank, 2 * world, world)
        refused = False
    except NotImplementedError:
        refused = True
    cu_q = torch.arange(B + 1, dtype=torch.int32, device=dev) * Ll
    cu_k = torch.arange(B + 1, dtype=torch.int32, device=dev) * L
    torch.manual_seed(4)
    qf = torch.randn(B, L, Hq, D, device=dev).to(dtype)
    kf = torch.randn(B, L, Hkv, D, device=dev).to(dtype)
    vf = torch.randn(B, L, Hkv, D, device=dev).to(dtype)
    pos = _global_pos(torch.arange(Ll, device=dev), L, rank, world, True)
    q, k, v = (t[:, pos].reshape(B * Ll, t.shape[2], D).contiguous() for t in (qf, kf, vf))
    out = fused_sp_ag_attn_inter_node(ctx, q, k, v, None, cu_q, cu_k, Ll, L, rank, world, world, True, True)
    g = Hq // Hkv
    ref = torch.nn.functional.scaled_dot_product_attention(qf.float().transpose(1, 2), kf.float().repeat_interleave(g, 2).transpose(1, 2),
                                                           vf.float().repeat_interleave(g, 2).transpose(1, 2), is_causal=True)
    ref = ref.transpose(1, 2)[:, pos].reshape(B * Ll, Hq, D)
    ok = all_ok(refused and check_close(out, ref, dtype if dev == "cuda" else torch.float16, "sp_ag_attn (inter-node API)"))
    ctx.finalize()
    finish(ok)
