"""GPU numerics of the decode-path kernels against plain PyTorch fp32 references."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,hidden", [(1, 5120), (7, 4096), (64, 1024)])
@pytest.mark.parametrize("with_res", [False, True])
def test_rmsnorm(rows, hidden, with_res):
    from triton_dist.kernels.nvidia.elementwise import rmsnorm

    torch.manual_seed(0)
    x = torch.randn(rows, hidden, device="cuda").bfloat16()
    w = (1 + 0.1 * torch.randn(hidden, device="cuda")).bfloat16()
    res = torch.randn(rows, hidden, device="cuda").bfloat16() if with_res else None
    res_ref = None if res is None else (x.float() + res.float()).bfloat16()
    xs = x.float() if res is None else res_ref.float()
    ref = xs * torch.rsqrt(xs.pow(2).mean(-1, keepdim=True) + 1e-6) * w.float()
    y = rmsnorm(x, w, 1e-6, residual=res)
    torch.testing.assert_close(y.float(), ref, atol=3e-2, rtol=2e-2)
    if with_res:
        torch.testing.assert_close(res.float(), res_ref.float(), atol=0, rtol=0)


def test_silu_mul_and_add():
    from triton_dist.kernels.nvidia.elementwise import add, silu_mul

    torch.manual_seed(0)
    gu = torch.randn(5, 2 * 3200, device="cuda").bfloat16()
    ref = torch.nn.functional.silu(gu[:, :3200].float()) * gu[:, 3200:].float()
    torch.testing.assert_close(silu_mul(gu).float(), ref, atol=3e-2, rtol=2e-2)
    a, b = torch.randn(3, 5120, device="cuda").bfloat16(), torch.randn(3, 5120, device="cuda").bfloat16()
    torch.testing.assert_close(add(a, b).float(), (a.float() + b.float()).bfloat16().float(), atol=0, rtol=0)


@pytest.mark.parametrize("M", [1, 2, 3, 8, 16])
def test_gemv(M):
    from triton_dist.kernels.nvidia.elementwise import gemv

    torch.manual_seed(0)
    x = torch.randn(M, 5120, device="cuda").bfloat16()
    w = (0.05 * torch.randn(3200, 5120, device="cuda")).bfloat16()
    y = gemv(x, w)
    torch.testing.assert_close(y.float(), x.float() @ w.float().t(), atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("paged", [False, True])
def test_qk_norm_rope_kvcache(paged):
    from triton_dist.kernels.nvidia.elementwise import make_cos_sin_cache, qk_norm_rope_update_kvcache, torch_qk_norm_rope

    torch.manual_seed(0)
    T, Hq, Hkv, D, S = 6, 8, 2, 128, 64
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device="cuda").bfloat16()
    qw = (1 + 0.1 * torch.randn(D, device="cuda")).bfloat16()
    kw = (1 + 0.1 * torch.randn(D, device="cuda")).bfloat16()
    cs = make_cos_sin_cache(D, S, device="cuda")
    pos = torch.tensor([3, 4, 5, 10, 0, 63], dtype=torch.int32, device="cuda")
    bidx = torch.tensor([0, 0, 0, 1, 2, 2], dtype=torch.int32, device="cuda")
    if paged:
        page = 16
        kc = torch.zeros(3 * S // page, page, Hkv, D, device="cuda", dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        bt = torch.randperm(3 * S // page, device="cuda").int().view(3, S // page)
    else:
        kc = torch.zeros(3, S, Hkv, D, device="cuda", dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        bt = None
    q = qk_norm_rope_update_kvcache(qkv, qw, kw, cs, pos, kc, vc, Hq, Hkv, D, 1e-6, batch_idx=bidx, block_table=bt)
    q_ref, k_ref, v_ref = torch_qk_norm_rope(qkv, qw, kw, cs, pos, Hq, Hkv, D, 1e-6)
    torch.testing.assert_close(q.float(), q_ref.float(), atol=3e-2, rtol=2e-2)
    for t in range(T):
        b, p = int(bidx[t]), int(pos[t])
        if paged:
            pg = int(bt[b, p // 16])
            kk, vv = kc[pg, p % 16], vc[pg, p % 16]
        else:
            kk, vv = kc[b, p], vc[b, p]
        torch.testing.assert_close(kk.float(), k_ref[t].float(), atol=3e-2, rtol=2e-2)
        torch.testing.assert_close(vv.float(), v_ref[t].float(), atol=0, rtol=0)


@pytest.mark.parametrize("B,Hq,Hkv,kv_len", [(1, 8, 1, 512), (2, 32, 8, 1000), (3, 16, 16, 77), (1, 64, 4, 4096)])
def test_flash_decode(B, Hq, Hkv, kv_len):
    from triton_dist.kernels.nvidia.flash_decode import (gqa_fwd_batch_decode, gqa_fwd_batch_decode_intra_rank,
                                                         inter_rank_combine, torch_paged_decode)

    torch.manual_seed(0)
    D, page = 128, 16
    pages_per_seq = (kv_len + page - 1) // page
    q = torch.randn(B, Hq, D, device="cuda").bfloat16()
    kc = torch.randn(B * pages_per_seq, page, Hkv, D, device="cuda").bfloat16()
    vc = torch.randn(B * pages_per_seq, page, Hkv, D, device="cuda").bfloat16()
    bt = torch.randperm(B * pages_per_seq, device="cuda").int().view(B, pages_per_seq)
    lens = torch.tensor([kv_len - 3 * i for i in range(B)], dtype=torch.int32, device="cuda")
    scale = 1.0 / math.sqrt(D)
    ref, lse_ref = torch_paged_decode(q, kc, vc, lens, bt, scale)
    out = gqa_fwd_batch_decode(q, kc, vc, None, [1] * B, lens, bt, scale)
    torch.testing.assert_close(out.float(), ref, atol=2e-2, rtol=2e-2)
    part = gqa_fwd_batch_decode_intra_rank(q, kc, vc, None, [1] * B, lens, bt, scale, kv_split=8)
    torch.testing.assert_close(part[..., :D], ref, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(part[..., D], lse_ref, atol=2e-2, rtol=1e-3)
    # inter-rank combine of two halves of the KV equals attention over the whole KV
    half = (lens // 2).int()
    # emulate two "ranks": rank 0 sees the first half of every sequence, rank 1 the whole (different lengths)
    p0 = gqa_fwd_batch_decode_intra_rank(q, kc, vc, None, [1] * B, half, bt, scale)
    # second shard: positions [half, len) -> build shifted block tables by re-packing the KV
    outs = []
    for b in range(B):
        n, h = int(lens[b]), int(half[b])
        pages = bt[b, :pages_per_seq].long()
        kf = kc[pages].reshape(-1, Hkv, D)[h:n]
        vf = vc[pages].reshape(-1, Hkv, D)[h:n]
        outs.append((kf, vf))
    m = max(o[0].shape[0] for o in outs)
    pps2 = (m + page - 1) // page
    kc2 = torch.zeros(B * pps2, page, Hkv, D, device="cuda", dtype=torch.bfloat16)
    vc2 = torch.zeros_like(kc2)
    bt2 = torch.arange(B * pps2, device="cuda", dtype=torch.int32).view(B, pps2)
    for b, (kf, vf) in enumerate(outs):
        kc2.view(B, pps2 * page, Hkv, D)[b, :kf.shape[0]] = kf
        vc2.view(B, pps2 * page, Hkv, D)[b, :vf.shape[0]] = vf
    p1 = gqa_fwd_batch_decode_intra_rank(q, kc2, vc2, None, [1] * B, (lens - half).int(), bt2, scale)
    merged = inter_rank_combine(torch.stack([p0, p1]))
    torch.testing.assert_close(merged.float(), ref, atol=2e-2, rtol=2e-2)


def test_embedding_argmax_linear():
    from triton_dist.kernels.nvidia.elementwise import argmax, embedding, linear

    torch.manual_seed(0)
    table = torch.randn(1000, 512, device="cuda").bfloat16()
    ids = torch.randint(0, 1000, (2, 5), device="cuda")
    torch.testing.assert_close(embedding(ids, table), table[ids])
    logits = torch.randn(3, 151936, device="cuda").bfloat16()
    assert torch.equal(argmax(logits), logits.float().argmax(-1))
    x = torch.randn(40, 1024, device="cuda").bfloat16()
    w = (0.05 * torch.randn(2048, 1024, device="cuda")).bfloat16()
    torch.testing.assert_close(linear(x, w).float(), x.float() @ w.float().t(), atol=3e-2, rtol=2e-2)
