"""GPU tests: numerics of every native kernel against a plain PyTorch fp32 reference.

Single-process tests call the C ABI through the Python wrappers; distributed ops run the
script-style tests under torchrun at world sizes 1 and min(#GPUs, 8)."""
import pytest
import torch

from conftest import gpu_count, run_dist

pytestmark = pytest.mark.gpu

N = "triton_dist.test.nvidia."


def _worlds():
    n = gpu_count()
    return sorted({1, min(n, 2), n} - {0}) if n else [1]


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
def test_gemm_every_tile_width(bn):
    from triton_dist.kernels.nvidia import gemm_persistent

    torch.manual_seed(1)
    a = (torch.rand(512, 1024, device="cuda") * 2 - 1).bfloat16()
    b = ((torch.rand(1344, 1024, device="cuda") * 2 - 1) * 0.1).bfloat16()
    c = gemm_persistent(a, b, block_n=bn)
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
def test_ag_gemm(world):
    run_dist(N + "test_ag_gemm", [2048 * world, 512 * world, 1024, "--check", "--stress"], world)


@pytest.mark.parametrize("world", _worlds())
def test_ag_gemm_ragged(world):
    run_dist(N + "test_ag_gemm", [200 * world, 136 * world, 520, "--check"], world)


@pytest.mark.parametrize("world", _worlds())
def test_gemm_rs(world):
    run_dist(N + "test_gemm_rs", [2048 * world, 1024, 512 * world, "--check", "--stress"], world)


@pytest.mark.parametrize("world", _worlds())
def test_gemm_ar(world):
    run_dist(N + "test_gemm_ar", [64, 1024, 512 * world, "--check", "--low-latency", "--stress"], world)
    run_dist(N + "test_gemm_ar", [2048, 1024, 256 * world, "--check"], world)


@pytest.mark.parametrize("world", _worlds())
def test_allreduce(world):
    run_dist(N + "test_allreduce", ["--check", "--max-bytes", 4 * 2**20, "--verify-iters", 2, "--stress"], world)


@pytest.mark.parametrize("world", _worlds())
def test_fast_allgather(world):
    run_dist(N + "test_fast_allgather", ["--check", "--max-bytes", 2**20, "--verify-iters", 2, "--stress"], world)


@pytest.mark.parametrize("world", _worlds())
def test_reduce_scatter(world):
    run_dist(N + "test_reduce_scatter", [1024 * world, 1024, "--check", "--stress"], world)
