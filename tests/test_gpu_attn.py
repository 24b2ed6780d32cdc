"""tcgen05 flash-attention forward (csrc/kernels/sp_attn_sm100.cu) against plain PyTorch fp32 attention: varlen, GQA,
causal / full, chunked prefill (queries = last rows of the sequence), dense-cache addressing, and the sequence-parallel
position mapping (plain and zig-zag) emulated on one GPU by looping over the ranks."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(q, k, v, cu_q, cu_k, causal, k_starts=None):
    """fp32: queries of sequence b are the last Lq positions of its L keys."""
    Hq, D = q.shape[1], q.shape[2]
    g = Hq // k.shape[1]
    out = torch.empty_like(q, dtype=torch.float32)
    for b in range(cu_q.numel() - 1):
        qs, qe = int(cu_q[b]), int(cu_q[b + 1])
        L = int(cu_k[b + 1] - cu_k[b])
        ks = int(cu_k[b]) if k_starts is None else k_starts[b]
        qq = q[qs:qe].float().transpose(0, 1)
        kk = k[ks:ks + L].float().repeat_interleave(g, dim=1).transpose(0, 1)
        vv = v[ks:ks + L].float().repeat_interleave(g, dim=1).transpose(0, 1)
        s = torch.einsum("hqd,hkd->hqk", qq, kk) / math.sqrt(D)
        if causal:
            pos = torch.arange(qe - qs, device=q.device) + (L - (qe - qs))
            s = s.masked_fill(~(torch.arange(L, device=q.device)[None, :] <= pos[:, None])[None], float("-inf"))
        out[qs:qe] = torch.einsum("hqk,hkd->hqd", torch.softmax(s, -1), vv).transpose(0, 1)
    return out


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("lens_q,lens_k,Hq,Hkv", [([128], [128], 4, 4), ([300, 77, 513], [300, 77, 513], 8, 2),
                                                   ([64, 200], [500, 1000], 4, 1), ([1024], [1024], 2, 2)])
def test_flash_attn_varlen_matches_fp32(causal, lens_q, lens_k, Hq, Hkv):
    from triton_dist.kernels.nvidia.sp_ag_attention_intra_node import flash_attn_varlen

    torch.manual_seed(0)
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32, device="cuda")
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32, device="cuda")
    q = torch.randn(sum(lens_q), Hq, 128, device="cuda").bfloat16()
    k = torch.randn(sum(lens_k), Hkv, 128, device="cuda").bfloat16()
    v = torch.randn(sum(lens_k), Hkv, 128, device="cuda").bfloat16()
    out = flash_attn_varlen(q, k, v, cu_q, cu_k, max(lens_q), causal=causal)
    torch.testing.assert_close(out.float(), _ref(q, k, v, cu_q, cu_k, causal), atol=2e-2, rtol=2e-2)


def test_flash_attn_dense_cache_with_garbage_behind_the_sequences():
    """k / v are a [B, S] cache whose rows behind each sequence hold NaN: they must not leak into the result."""
    from triton_dist.kernels.nvidia.sp_ag_attention_intra_node import flash_attn_varlen

    torch.manual_seed(1)
    B, S, Hq, Hkv, lens = 3, 640, 8, 4, [640, 129, 333]
    kc = torch.full((B, S, Hkv, 128), float("nan"), device="cuda", dtype=torch.bfloat16)
    vc = torch.full((B, S, Hkv, 128), float("nan"), device="cuda", dtype=torch.bfloat16)
    for b, n in enumerate(lens):
        kc[b, :n] = torch.randn(n, Hkv, 128, device="cuda").bfloat16()
        vc[b, :n] = torch.randn(n, Hkv, 128, device="cuda").bfloat16()
    q = torch.randn(sum(lens), Hq, 128, device="cuda").bfloat16()
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
    out = flash_attn_varlen(q, kc.view(B * S, Hkv, 128), vc.view(B * S, Hkv, 128), cu, cu, max(lens), causal=True, k_batch_stride=S)
    ref = _ref(q, kc.view(B * S, Hkv, 128), vc.view(B * S, Hkv, 128), cu, cu, True, k_starts=[b * S for b in range(B)])
    assert bool(torch.isfinite(out.float()).all())
    torch.testing.assert_close(out.float(), ref, atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("zigzag", [True, False])
def test_flash_attn_sequence_parallel_positions(zigzag):
    """Every rank's share of the queries against the gathered K/V, with the rank-dependent global positions."""
    from triton_dist.kernels.nvidia.sp_ag_attention_intra_node import flash_attn_varlen, torch_sp_attention

    torch.manual_seed(2)
    W, Hq, Hkv = 4, 4, 2
    lens = [512, 1024]  # divisible by 2 W
    cu_k = torch.tensor([0, 512, 1536], dtype=torch.int32, device="cuda")
    cu_q = (cu_k // W).to(torch.int32)
    k = torch.randn(1536, Hkv, 128, device="cuda").bfloat16()
    v = torch.randn(1536, Hkv, 128, device="cuda").bfloat16()
    for rank in range(W):
        q = torch.randn(1536 // W, Hq, 128, device="cuda").bfloat16()
        out = flash_attn_varlen(q, k, v, cu_q, cu_k, max(lens) // W, causal=True, rank=rank, world=W, zigzag=zigzag)
        ref = torch_sp_attention(q, k, v, cu_q, cu_k, rank, W, True, zigzag)
        torch.testing.assert_close(out.float(), ref, atol=2e-2, rtol=2e-2)
