"""GPU tests of the MoE / EP kernels and of the TP models (native kernels vs torch oracles / HF)."""
import pytest
import torch

from conftest import gpu_count, run_dist

pytestmark = pytest.mark.gpu
N = "triton_dist.test.nvidia."


def _worlds():
    n = gpu_count()
    return sorted({1, n} - {0}) if n else [1]


@pytest.mark.parametrize("T,E,topk", [(37, 8, 2), (1024, 128, 8), (4096, 256, 8), (5, 64, 4)])
def test_moe_align_sort_matches_oracle(T, E, topk):
    from triton_dist.kernels.nvidia.moe_utils import moe_align_sort, torch_align_sort

    torch.manual_seed(0)
    ids = torch.rand(T, E, device="cuda").topk(topk, dim=-1).indices.int()
    got, ref = moe_align_sort(ids, E), torch_align_sort(ids, E)
    assert torch.equal(got.meta, ref.meta)
    assert torch.equal(got.block_expert, ref.block_expert)
    # same multiset of pairs inside every expert's row range (order inside an expert is free)
    flat = ids.reshape(-1)
    rows = got.pair_to_row.long()
    assert torch.equal(got.sorted_pair_ids[rows].long(), torch.arange(T * topk, device="cuda"))
    assert torch.equal(got.block_expert[rows // 128].long(), flat.long())
    assert int((got.sorted_pair_ids < T * topk).sum()) == T * topk


@pytest.mark.parametrize("T,H,I,E,topk", [(256, 1024, 512, 16, 4), (77, 512, 256, 8, 2)])
def test_grouped_gemm_pipeline(T, H, I, E, topk):
    from triton_dist.kernels.nvidia.elementwise import silu_mul
    from triton_dist.kernels.nvidia.moe_utils import grouped_gemm, moe_align_sort, moe_gather_rows, reduce_topk

    torch.manual_seed(1)
    x = torch.randn(T, H, device="cuda").bfloat16()
    w1 = (0.05 * torch.randn(E, 2 * I, H, device="cuda")).bfloat16()
    w2 = (0.05 * torch.randn(E, H, I, device="cuda")).bfloat16()
    probs = torch.softmax(torch.randn(T, E, device="cuda"), -1)
    wt, ids = probs.topk(topk, dim=-1)
    sort = moe_align_sort(ids, E)
    h = silu_mul(grouped_gemm(moe_gather_rows(x, sort), w1, sort))
    out = reduce_topk(grouped_gemm(h, w2, sort), sort, wt)
    ref = torch.zeros(T, H, device="cuda")
    for e in range(E):
        tok, k = (ids == e).nonzero(as_tuple=True)
        if tok.numel():
            g, u = (x[tok].float() @ w1[e].float().t()).bfloat16().float().chunk(2, -1)
            hh = (torch.nn.functional.silu(g) * u).bfloat16().float()
            ref.index_add_(0, tok, (hh @ w2[e].float().t()).bfloat16().float() * wt[tok, k, None])
    torch.testing.assert_close(out.float(), ref, atol=3e-2 * float(ref.abs().max()), rtol=3e-2)


@pytest.mark.parametrize("world", _worlds())
def test_distributed_moe_suite(world):
    """EP dispatch/combine (bf16 + fp8), low-latency all-to-all, TP_MoE (AG+MoE / MoE+RS), distributed flash-decode."""
    run_dist("triton_dist.test.nvidia.run_suite", ["--suite", "moe"], world, timeout=900)


@pytest.mark.parametrize("world", _worlds())
def test_distributed_models_suite(world):
    """DenseLLM / Qwen3MoE in every backend mode vs HuggingFace, and the decode megakernel (fused + unfused graph)."""
    run_dist("triton_dist.test.nvidia.run_suite", ["--suite", "models"], world, timeout=900)


@pytest.mark.gpu
def test_moe_tile_swizzle_native_vs_torch():
    """AG+MoE block schedule: the native single-CTA kernel against the host restatement (reference K45)."""
    import torch

    from triton_dist.kernels.nvidia.moe_utils import moe_align_sort
    from triton_dist.kernels.nvidia.threadblock_swizzle_ag_moe import (threadblock_swizzle_ag_moe,
                                                                       torch_threadblock_swizzle_ag_moe)

    torch.manual_seed(0)
    ids = (torch.randn(1024, 32, device="cuda") + torch.linspace(-1, 1, 32, device="cuda")).topk(6, dim=-1).indices
    s = moe_align_sort(ids, 32)
    for rank in (0, 3):
        got = threadblock_swizzle_ag_moe(s, 256, rank, 4)
        ref = torch_threadblock_swizzle_ag_moe(s, 256, rank, 4)
        nb = int(s.meta[1])
        assert torch.equal(got.segment_start[:nb].cpu(), ref.segment_start.cpu())
        assert torch.equal(got.segment_end[:nb].cpu(), ref.segment_end.cpu())
        assert torch.equal(got.block_order[:nb].cpu(), ref.block_order.cpu())
        assert torch.equal(got.stage_count.cpu(), ref.stage_count.cpu())
