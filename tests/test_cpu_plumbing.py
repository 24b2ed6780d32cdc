"""BASELINE config 1: the op API end to end on CPU/gloo, world_size 2 (no GPU needed).

The fused ops take their torch.distributed path here; what is exercised is the
context/op surface, shape logic and the multi-process launcher.

Every case is one torchrun launch (a few seconds of work behind ~10 s of interpreter start-up per rank), so the launches of
this file are started ahead by a small pool when its first test runs; the tests then look at the results."""
import os
from concurrent.futures import ThreadPoolExecutor

import pytest

from conftest import run_dist

N = "triton_dist.test.nvidia."
O = "triton_dist.mega_triton_kernel.test.ops."
_MOE_SMALL = ["-M", 512, "-E", 8, "--topk", 2]
# name: (module, arguments, ranks, timeout in seconds)
JOBS = {
    "ag_gemm[1]": (N + "test_ag_gemm", [1024, 1024, 1024, "--check", "--verify-iters", 2], 1, 600),
    "ag_gemm[2]": (N + "test_ag_gemm", [1024, 1024, 1024, "--check", "--verify-iters", 2], 2, 600),
    "gemm_rs[2]": (N + "test_gemm_rs", [1024, 1024, 1024, "--check", "--verify-iters", 2], 2, 600),
    "gemm_ar": (N + "test_gemm_ar", [64, 512, 1024, "--check", "--verify-iters", 2], 2, 600),
    "allreduce": (N + "test_allreduce", ["--check", "--max-bytes", 65536, "--verify-iters", 1, "--method", "auto"], 2, 600),
    "fast_allgather": (N + "test_fast_allgather", ["--check", "--max-bytes", 16384, "--verify-iters", 1, "--mode", "pull"], 2, 600),
    "reduce_scatter": (N + "test_reduce_scatter", [256, 128, "--check", "--verify-iters", 1], 2, 600),
    "tp_e2e[qwen3]": (N + "test_tp_e2e", ["--model", "qwen3", "--check"], 2, 900),
    "tp_e2e[llama]": (N + "test_tp_e2e", ["--model", "llama", "--check"], 2, 900),
    "tp_e2e[qwen3_moe]": (N + "test_tp_e2e", ["--model", "qwen3_moe", "--check"], 2, 900),
    "tp_e2e_corrupt": (N + "test_tp_e2e", ["--model", "qwen3", "--check", "--corrupt", "--modes", "torch,triton_dist_AR"], 2, 900),
    "ep_a2a": (N + "test_ep_a2a", ["--check", "--verify-iters", 2], 2, 600),
    "all_to_all": (N + "test_all_to_all", ["--check", "--verify-iters", 2, "--with-scale"], 2, 600),
    "tp_moe": (N + "test_tp_moe", ["--check", "--verify-iters", 1, "-M", 64, "--hidden", 64, "--inter", 32, "-E", 4, "--topk", 2], 2, 600),
    "ag_moe_overlap[2]": (N + "test_ag_moe", ["--check", "--verify-iters", 2, "--overlap"] + _MOE_SMALL + ["--hidden", 256, "--inter", 512], 2, 600),
    "ag_moe_overlap[4]": (N + "test_ag_moe", ["--check", "--verify-iters", 2, "--overlap"] + _MOE_SMALL + ["--hidden", 256, "--inter", 512], 4, 600),
    "moe_rs_overlap[2]": (N + "test_moe_reduce_rs", ["--check", "--verify-iters", 2, "--overlap"] + _MOE_SMALL + ["--hidden", 1024, "--inter", 256], 2, 600),
    "moe_rs_overlap[4]": (N + "test_moe_reduce_rs", ["--check", "--verify-iters", 2, "--overlap"] + _MOE_SMALL + ["--hidden", 1024, "--inter", 256], 4, 600),
    "moe_overlap_single": (N + "test_moe_overlap_single_gpu", [], 1, 600),
    "sp_decode": (N + "test_sp_decode_attn", ["--check", "--verify-iters", 1, "--kv-len", 256, "--q-heads", 4, "--kv-heads", 2], 2, 600),
    "suite[layers]": (N + "run_suite", ["--suite", "layers"], 2, 1200),
    "suite[prims]": (N + "run_suite", ["--suite", "prims"], 2, 600),
    "suite[mega]": (N + "run_suite", ["--suite", "mega"], 2, 1200),
    "team_split[2]": (N + "test_team_split", [], 2, 600),
    "team_split[4]": (N + "test_team_split", [], 4, 600),
}
JOBS.update({f"mega_op[{n}]": (O + n, [], 2, 600) for n in ("test_add", "test_rms_norm", "test_mlp_layer", "test_allreduce", "test_page_attn")})
_pool, _futures = None, {}


def _launch(module, args, nproc, timeout):
    try:
        return run_dist(module, args, nproc, timeout=timeout)
    except AssertionError as e:
        if "address already in use" not in str(e).lower() and "eaddrinuse" not in str(e).lower():
            raise
        return run_dist(module, args, nproc, timeout=timeout)  # two launches picked the same free port: once more


def job(name: str) -> str:
    """Output of the named launch (an AssertionError of the launcher is re-raised here, in the test that asked)."""
    global _pool
    if _pool is None:
        cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 2)
        _pool = ThreadPoolExecutor(max_workers=max(1, min(4, cores // 2)))  # a launch is 2-4 ranks, mostly interpreter start-up
        for k, spec in JOBS.items():
            _futures[k] = _pool.submit(_launch, *spec)
    if _futures[name].cancel():  # still queued (e.g. a run with -k): run it here instead of waiting for its turn
        return _launch(*JOBS[name])
    return _futures[name].result()


@pytest.fixture(scope="module", autouse=True)
def _drop_unasked_launches():
    yield
    for f in _futures.values():  # (a run with -k: do not finish the rest of the queue at exit)
        f.cancel()


@pytest.mark.parametrize("world", [1, 2])
def test_ag_gemm_cpu(world):
    job(f"ag_gemm[{world}]")


@pytest.mark.parametrize("world", [2])
def test_gemm_rs_cpu(world):
    job(f"gemm_rs[{world}]")


def test_gemm_ar_cpu():
    job("gemm_ar")


def test_allreduce_cpu():
    job("allreduce")


def test_fast_allgather_cpu():
    job("fast_allgather")


def test_reduce_scatter_cpu():
    job("reduce_scatter")


@pytest.mark.parametrize("model", ["qwen3", "llama", "qwen3_moe"])
def test_tp_e2e_vs_hf_cpu(model):
    """DenseLLM / Qwen3MoE sharded over 2 ranks reproduce the HF model's logits in every backend mode."""
    job(f"tp_e2e[{model}]")


def test_tp_e2e_comparison_can_fail_cpu():
    """The HF comparison is not vacuous: with ten LM-head rows swapped every backend mode must be reported as a mismatch."""
    job("tp_e2e_corrupt")


def test_ep_dispatch_combine_cpu():
    job("ep_a2a")


def test_low_latency_all_to_all_cpu():
    job("all_to_all")


def test_tp_moe_cpu():
    job("tp_moe")


@pytest.mark.parametrize("world", [2, 4])
def test_moe_overlap_schedules_cpu(world):
    """Stage-gated AG + grouped GEMM (per-stage block masks, stage-filtered gather) and the column-chunked MoE + RS give the
    same results as the per-expert torch loop: the orchestration of the opt-in overlap schedules, on their CPU paths."""
    job(f"ag_moe_overlap[{world}]")
    job(f"moe_rs_overlap[{world}]")


def test_moe_overlap_single_device_cpu():
    """Stage-gated schedule as every rank of a pretend world of 4 (masks partition the blocks, results bit-identical to the
    one-shot form) and the chunked MoE + RS pipeline at world size 1."""
    job("moe_overlap_single")


def test_sp_flash_decode_cpu():
    job("sp_decode")


def test_layers_suite_cpu():
    """TP_MLP / TP_Attn / MoE / EP-MoE / decode attention / Engine.serve scripts on their torch paths (gloo, 2 ranks)."""
    job("suite[layers]")


def test_prims_suite_cpu():
    """The primitive / host-API scripts (ring put, notify, waits, heap API under its NVSHMEM names, host utils, small-message
    all-gathers, PP) on the host heap: the parts that need CUDA peer mappings skip themselves, the rest must pass."""
    job("suite[prims]")


def test_mega_suite_cpu():
    """ModelBuilder graphs (per-op cases and the Qwen3 decode graph) through the eager fallback of the builder."""
    job("suite[mega]")


@pytest.mark.parametrize("name", ["test_add", "test_rms_norm", "test_mlp_layer", "test_allreduce", "test_page_attn"])
def test_mega_per_op_scripts_cpu(name):
    """The per-op megakernel scripts (same file names as the reference's mega_triton_kernel/test/ops): their oracles agree
    with the builder's eager fallback, so on a GPU any mismatch points at the task body."""
    job(f"mega_op[{name}]")


@pytest.mark.parametrize("world", [2, 4])
def test_team_split_cpu(world):
    """EP x PP team split, rank translation, strided split (reference test_team_split.py)."""
    job(f"team_split[{world}]")


def test_model_server_and_chat_cpu():
    """model_server (2 ranks, random-init Qwen3-0.6B on the builder's eager path) answers a request sent by the chat client:
    prompt broadcast to all ranks, token-by-token prefill + greedy decode, JSON reply (reference model_server.py / chat.py)."""
    import os
    import subprocess
    import sys
    import time

    from conftest import ROOT, free_port

    port = free_port()
    env = dict(os.environ)
    env["PYTHONPATH"] = str(ROOT) + os.pathsep + env.get("PYTHONPATH", "")
    env["RANDOM_PARAMS"] = "1"
    env.setdefault("OMP_NUM_THREADS", "2")
    server = subprocess.Popen([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                               "127.0.0.1", "--master-port", str(free_port()), "-m",
                               "triton_dist.mega_triton_kernel.test.models.model_server", "--model", "Qwen/Qwen3-0.6B", "--port", str(port),
                               "--max-length", "64", "--max-requests", "1"], cwd=str(ROOT), env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True)
    try:
        deadline = time.time() + 600
        for line in server.stdout:  # wait until rank 0 listens
            if "model_server listening" in line:
                break
            assert time.time() < deadline, "server did not come up"
        else:
            raise AssertionError("server exited before listening")
        client = subprocess.run([sys.executable, "-m", "triton_dist.mega_triton_kernel.test.models.chat", "--port", str(port),
                                 "--prompt-ids", "11,22,33", "--max-new-tokens", "4"], cwd=str(ROOT), env=env, capture_output=True,
                                text=True, timeout=600)
        assert client.returncode == 0, client.stdout + client.stderr
        ids = eval(client.stdout.strip().splitlines()[-1])  # the client prints the list of generated token ids
        assert isinstance(ids, list) and len(ids) == 4 and all(isinstance(t, int) and 0 <= t < 151936 for t in ids), client.stdout
        assert server.wait(timeout=300) == 0
    finally:
        if server.poll() is None:
            server.kill()
