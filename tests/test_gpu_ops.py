"""GPU tests: numerics of every native kernel against a plain PyTorch fp32 reference.

Single-process tests call the C ABI through the Python wrappers; distributed ops run as suites of the script-style
tests (one torchrun launch per suite and world size: 1 GPU and all GPUs of the box)."""
import pytest
import torch

from conftest import gpu_count, run_dist

pytestmark = pytest.mark.gpu


def _worlds():
    n = gpu_count()
    return sorted({1, n} - {0}) if n else [1]


def test_native_library_is_loaded():
    from triton_dist import _lib

    assert _lib.use_native()
    assert _lib.lib() is not None  # raises if libtdist.so is missing: no silent fallback on a GPU box


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 4096), (8192, 1792, 512), (1000, 1032, 520), (96, 136, 72),
                                   (4096, 14336, 1024)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_gemm_matches_fp32_reference(M, N, K, dtype):
    from triton_dist.kernels.nvidia import gemm_persistent

    torch.manual_seed(0)
    a = (torch.rand(M, K, device="cuda") * 2 - 1).to(dtype)
    b = ((torch.rand(N, K, device="cuda") * 2 - 1) * 0.1).to(dtype)
    c = gemm_persistent(a, b)
    ref = a.float() @ b.float().t()
    torch.testing.assert_close(c.float(), ref, atol=2e-2 if dtype == torch.bfloat16 else 5e-3, rtol=2e-2)


@pytest.mark.parametrize("bn", [32, 64, 96, 112, 128, 160, 192, 224, 256])
@pytest.mark.parametrize("cta_group", [1, 2])
def test_gemm_every_tile_width(bn, cta_group):
    """Every tile-width instantiation of the single-CTA kernel and of the CTA-pair kernel (which has no 32-wide tile: the
    request then falls back to its cost model)."""
    from triton_dist.kernels.nvidia import gemm_persistent

    torch.manual_seed(1)
    a = (torch.rand(512, 1024, device="cuda") * 2 - 1).bfloat16()
    b = ((torch.rand(1344, 1024, device="cuda") * 2 - 1) * 0.1).bfloat16()
    c = gemm_persistent(a, b, block_n=bn, cta_group=cta_group)
    torch.testing.assert_close(c.float(), a.float() @ b.float().t(), atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("M", [8, 16, 64])
def test_gemm_skinny_swap_ab(M):
    from triton_dist.kernels.nvidia import gemm_persistent

    torch.manual_seed(2)
    a = (torch.rand(M, 2048, device="cuda") * 2 - 1).bfloat16()
    b = ((torch.rand(5120, 2048, device="cuda") * 2 - 1) * 0.1).bfloat16()
    c = gemm_persistent(a, b)
    torch.testing.assert_close(c.float(), a.float() @ b.float().t(), atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("world", _worlds())
def test_distributed_core_suite(world):
    """AG+GEMM, GEMM+RS, GEMM+AR, AllReduce (all methods), fast AllGather (all modes), ReduceScatter."""
    run_dist("triton_dist.test.nvidia.run_suite", ["--suite", "core"], world, timeout=900)
