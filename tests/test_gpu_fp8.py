"""Block-scaled fp8 GEMM (tcgen05.mma.kind::f8f6f4, two-level accumulation) against a plain PyTorch fp32 reference of the
SAME quantised operands: the kernel multiplies exact e4m3 values and accumulates in fp32, so only the summation order and
the 16-bit output rounding differ."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a8, sa, b8, sb):
    from triton_dist.kernels.nvidia.gemm_fp8 import dequant_1x128, dequant_128x128

    return dequant_1x128(a8, sa) @ dequant_128x128(b8, sb).t()


@pytest.mark.parametrize("M,N,K", [(128, 128, 128), (256, 512, 1024), (1000, 1032, 528), (96, 136, 272), (4096, 2048, 7168)])
def test_gemm_fp8_matches_fp32_reference(M, N, K):
    from triton_dist.kernels.nvidia.gemm_fp8 import gemm_fp8, quant_fp8_1x128, quant_fp8_128x128

    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda") * (1 + 4 * torch.rand(M, 1, device="cuda"))  # rows of different magnitude
    b = torch.randn(N, K, device="cuda") * 0.05
    a8, sa = quant_fp8_1x128(a)
    b8, sb = quant_fp8_128x128(b)
    c = gemm_fp8(a8, sa, b8, sb)
    ref = _ref(a8, sa, b8, sb)
    torch.testing.assert_close(c.float(), ref, atol=2e-2 * float(ref.abs().max()), rtol=2e-2)
    # and the quantised product stays close to the unquantised one (sanity of scales / layout, loose)
    full = a @ b.t()
    assert float((c.float() - full).abs().max()) < 0.08 * float(full.abs().max())


def test_gemm_fp8_grouped_with_padding_blocks():
    from triton_dist.kernels.nvidia.gemm_fp8 import gemm_fp8, quant_fp8_1x128, quant_fp8_128x128

    torch.manual_seed(1)
    E, N, K, blocks = 4, 256, 512, 7
    a = torch.randn(blocks * 128, K, device="cuda")
    w = torch.randn(E, N, K, device="cuda") * 0.05
    block_expert = torch.tensor([2, 0, -1, 3, 3, 1, -1], dtype=torch.int32, device="cuda")
    a8, sa = quant_fp8_1x128(a)
    w8, sw = quant_fp8_128x128(w)
    out = torch.full((blocks * 128, N), 7.0, device="cuda", dtype=torch.bfloat16)
    gemm_fp8(a8, sa, w8, sw, out=out, block_expert=block_expert)
    for blk, e in enumerate(block_expert.tolist()):
        rows = slice(blk * 128, (blk + 1) * 128)
        if e < 0:
            assert bool((out[rows] == 7.0).all())  # padding blocks are not touched
        else:
            ref = _ref(a8[rows], sa[rows], w8[e], sw[e])
            torch.testing.assert_close(out[rows].float(), ref, atol=2e-2 * float(ref.abs().max()), rtol=2e-2)
