"""Megakernel attention tasks against a dense torch oracle (counterpart of the reference's
mega_triton_kernel/test/ops/test_page_attn.py): q/k-norm + RoPE + paged KV append (``make_qk_norm_rope_update_kvcache``),
split-KV paged GQA decode + LSE combine (``make_attn``), and the fused single-task form (``make_attn_fused``).
The oracle gathers the pages into a dense [len, Hkv, D] K/V, applies its own norm / rotation and a plain softmax.
    python -m triton_dist.mega_triton_kernel.test.ops.test_page_attn"""
import torch

from triton_dist.kernels.nvidia.elementwise import make_cos_sin_cache
from triton_dist.mega_triton_kernel import ModelBuilder
from triton_dist.mega_triton_kernel.test.ops.test_ops import check
from triton_dist.test.common import all_ok, finish, setup
from triton_dist.utils import dist_print

D = 128


def _norm(t, w, eps):
    return t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + eps) * w.float()


def _rope(t, cos_sin, pos):  # t [H, D] fp32, rotate-half convention, cache rows = [cos(D/2) | sin(D/2)]
    half = D // 2
    cos, sin = cos_sin[pos, :half], cos_sin[pos, half:]
    return torch.cat([t[:, :half] * cos - t[:, half:] * sin, t[:, half:] * cos + t[:, :half] * sin], dim=-1)


def attn_case(rank, world, B, Hq, Hkv, ctx_lens, page, fused, dev, dtype, eps=1e-6):
    torch.manual_seed(41 + B + Hq + sum(ctx_lens))
    max_pos = max(ctx_lens) + 8
    pages_per_seq = (max_pos + page - 1) // page
    n_pages = B * pages_per_seq
    k_cache = torch.randn(n_pages, page, Hkv, D, device=dev, dtype=dtype)
    v_cache = torch.randn(n_pages, page, Hkv, D, device=dev, dtype=dtype)
    # a scrambled page table: sequence b owns pages perm[b * pps : (b + 1) * pps]
    perm = torch.randperm(n_pages, device=dev).to(torch.int32)
    block_table = perm.view(B, pages_per_seq).contiguous()
    positions = torch.tensor(ctx_lens, device=dev, dtype=torch.int32)  # position written by this step = old length
    qkv = torch.randn(B, 1, Hq + 2 * Hkv, D, device=dev, dtype=dtype)
    qw = (torch.rand(D, device=dev) + 0.5).to(dtype)
    kw = (torch.rand(D, device=dev) + 0.5).to(dtype)
    cos_sin = make_cos_sin_cache(D, max_pos, 1e6, device=dev)
    q_rope = torch.zeros(B, 1, Hq, D, device=dev, dtype=dtype)
    out = torch.zeros(B, 1, Hq, D, device=dev, dtype=dtype)
    scale = D ** -0.5

    # ---- oracle on a copy of the caches (the graph appends K/V of the new token in place)
    kc, vc = k_cache.clone(), v_cache.clone()
    ref = torch.zeros(B, Hq, D, device=dev)
    ref_q = torch.zeros(B, Hq, D, device=dev)
    g = Hq // Hkv
    for b in range(B):
        pos = ctx_lens[b]
        x = qkv[b, 0].float()
        q = _rope(_norm(x[:Hq], qw, eps), cos_sin, pos).to(dtype).float()
        k_new = _rope(_norm(x[Hq:Hq + Hkv], kw, eps), cos_sin, pos).to(dtype)
        v_new = x[Hq + Hkv:].to(dtype)
        pg = int(block_table[b, pos // page])
        kc[pg, pos % page], vc[pg, pos % page] = k_new, v_new
        pages = block_table[b, :(pos + page) // page].long()
        K = kc[pages].reshape(-1, Hkv, D)[:pos + 1].float().repeat_interleave(g, dim=1)  # [n, Hq, D]
        V = vc[pages].reshape(-1, Hkv, D)[:pos + 1].float().repeat_interleave(g, dim=1)
        p = torch.softmax(torch.einsum("hd,nhd->hn", q, K) * scale, dim=-1)
        ref[b] = torch.einsum("hn,nhd->hd", p, V)
        ref_q[b] = q

    bld = ModelBuilder(rank, world, world)
    if fused:
        bld.make_attn_fused(qkv, k_cache, v_cache, block_table, positions, qw, kw, cos_sin, out, eps)
    else:
        bld.make_qk_norm_rope_update_kvcache(qkv, k_cache, v_cache, block_table, positions, qw, kw, cos_sin, q_rope, eps, eps)
        bld.make_attn(q_rope, k_cache, v_cache, block_table, positions, out)
    bld.compile()
    ok = True
    name = f"attn B={B} Hq={Hq} Hkv={Hkv} ctx={ctx_lens} page={page} fused={fused}"
    for it in range(2):  # the second launch rewrites the same position with the same values: results must not move
        out.zero_()
        bld.run()
        if dev == "cuda":
            torch.cuda.synchronize()
        ok = check(f"{name} run{it}", out.view(B, Hq, D), ref, 2e-2) and ok
        if not fused:
            ok = check(f"{name} q_rope run{it}", q_rope.view(B, Hq, D), ref_q, 2e-2) and ok
        for b in range(B):  # the new K/V row landed in the right page
            pos = ctx_lens[b]
            pg = int(block_table[b, pos // page])
            ok = check(f"{name} k append b={b}", k_cache[pg, pos % page], kc[pg, pos % page].float(), 2e-2) and ok
            ok = check(f"{name} v append b={b}", v_cache[pg, pos % page], vc[pg, pos % page].float(), 1e-6) and ok
    bld.finalize()
    return ok


if __name__ == "__main__":
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    ok = True
    # (B, Hq, Hkv, context lengths, page size, fused): one and several KV heads, lengths around page and split boundaries
    for case in [(1, 8, 1, [511], 16, False), (1, 8, 1, [511], 16, True), (2, 16, 2, [1, 700], 16, False),
                 (2, 16, 2, [1, 700], 16, True), (1, 64, 8, [2047], 32, False), (4, 8, 8, [0, 15, 16, 333], 16, True)]:
        ok = attn_case(rank, world, *case, dev, dtype) and ok
    ok = all_ok(ok)
    dist_print("[mega page_attn] " + ("PASS" if ok else "FAIL"))
    finish(ok)
