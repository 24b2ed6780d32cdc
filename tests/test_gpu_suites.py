"""Distributed suites added after the core / moe / models ones: megakernel per-op cases, per-layer scripts, primitives.
One torchrun launch per (suite, world size): 1 GPU and all GPUs of the box."""
import pytest

from conftest import gpu_count, run_dist

pytestmark = pytest.mark.gpu


def _worlds():
    n = gpu_count()
    return sorted({1, n} - {0}) if n else [1]


@pytest.mark.parametrize("world", _worlds())
def test_distributed_mega_suite(world):
    """Megakernel task bodies per op (linear shapes, chained graph, cross-rank all-reduce) and the Qwen3 graphs."""
    run_dist("triton_dist.test.nvidia.run_suite", ["--suite", "mega"], world, timeout=900)


@pytest.mark.parametrize("world", _worlds())
def test_distributed_layers_suite(world):
    """TP_MLP / TP_Attn in every mode, MoE utils, AG+MoE, MoE+RS, EP MoE layer (bf16 + fp8), decode attention, Engine.serve."""
    run_dist("triton_dist.test.nvidia.run_suite", ["--suite", "layers"], world, timeout=900)


@pytest.mark.parametrize("world", _worlds())
def test_distributed_prims_suite(world):
    """ring put, notify, waits, barriers, multimem, heap API, host utils, copy-engine and small-message all-gathers, PP."""
    run_dist("triton_dist.test.nvidia.run_suite", ["--suite", "prims"], world, timeout=900)
