"""BASELINE config 1: the op API end to end on CPU/gloo, world_size 2 (no GPU needed).

The fused ops take their torch.distributed path here; what is exercised is the
context/op surface, shape logic and the multi-process launcher."""
import pytest

from conftest import run_dist

N = "triton_dist.test.nvidia."


@pytest.mark.parametrize("world", [1, 2])
def test_ag_gemm_cpu(world):
    run_dist(N + "test_ag_gemm", [1024, 1024, 1024, "--check", "--verify-iters", 2], world)


@pytest.mark.parametrize("world", [2])
def test_gemm_rs_cpu(world):
    run_dist(N + "test_gemm_rs", [1024, 1024, 1024, "--check", "--verify-iters", 2], world)


def test_gemm_ar_cpu():
    run_dist(N + "test_gemm_ar", [64, 512, 1024, "--check", "--verify-iters", 2], 2)


def test_allreduce_cpu():
    run_dist(N + "test_allreduce", ["--check", "--max-bytes", 65536, "--verify-iters", 1, "--method", "auto"], 2)


def test_fast_allgather_cpu():
    run_dist(N + "test_fast_allgather", ["--check", "--max-bytes", 16384, "--verify-iters", 1, "--mode", "pull"], 2)


def test_reduce_scatter_cpu():
    run_dist(N + "test_reduce_scatter", [256, 128, "--check", "--verify-iters", 1], 2)
