"""Run a named group of distributed checks in one torchrun launch.
    torchrun --nproc-per-node N -m triton_dist.test.nvidia.run_suite --suite core|moe|models|mega|layers|prims|all"""
import argparse
import os
import sys

import torch

from triton_dist.test.common import run_suite

T = "triton_dist.test.nvidia."
M = "triton_dist.mega_triton_kernel.test.models."


def cases(suite: str, world: int):
    core = [
        (T + "test_ag_gemm", [2048 * world, 512 * world, 1024, "--check", "--stress", "--alternate"]),
        (T + "test_ag_gemm", [200 * world, 136 * world, 520, "--check", "--verify-iters", 2]),
        (T + "test_ag_gemm", [512 * world, 256 * world, 512, "--check", "--verify-iters", 1, "--autotune"]),
        (T + "test_gemm_rs", [512 * world, 512, 256 * world, "--check", "--verify-iters", 1, "--autotune"]),
        (T + "test_gemm_rs", [2048 * world, 1024, 512 * world, "--check", "--stress", "--alternate"]),
        (T + "test_gemm_rs", [200 * world, 136, 72 * world, "--check", "--verify-iters", 2]),
        (T + "test_gemm_ar", [64, 1024, 512 * world, "--check", "--low-latency", "--stress"]),
        (T + "test_gemm_ar", [2048, 1024, 256 * world, "--check", "--verify-iters", 2]),
        (T + "test_allreduce", ["--check", "--max-bytes", 4 * 2**20, "--verify-iters", 2, "--stress"]),
        (T + "test_fast_allgather", ["--check", "--max-bytes", 2**20, "--verify-iters", 2, "--stress"]),
        (T + "test_reduce_scatter", [1024 * world, 1024, "--check", "--stress"]),
    ]
    moe = [
        (T + "test_ep_a2a", ["--check", "--verify-iters", 3, "--stress"]),
        (T + "test_ep_a2a", ["--check", "--verify-iters", 2, "--fp8"]),
        (T + "test_all_to_all", ["--check", "--verify-iters", 3, "--with-scale", "--stress"]),
        (T + "test_tp_moe", ["--check", "--verify-iters", 2, "--stress"]),
        (T + "test_sp_decode_attn", ["--check", "--verify-iters", 2, "--stress"]),
        (T + "test_sp_ag_attention_intra_node", ["--check", "--verify-iters", 2, "--seqlen", 512 * max(world, 1)]),
        (T + "test_sp_ag_attention_intra_node", ["--check", "--verify-iters", 1, "--seqlen", 256 * max(world, 1), "--no-zigzag"]),
        (T + "test_sp_ag_attention_inter_node", ["--check"]),
    ]
    models = [
        (T + "test_tp_e2e", ["--model", "qwen3", "--check"]),
        (T + "test_tp_e2e", ["--model", "llama", "--check"]),
        (T + "test_tp_e2e", ["--model", "qwen3_moe", "--check"]),
        (T + "test_tp_e2e", ["--model", "qwen3", "--check", "--corrupt", "--modes", "torch,triton_dist_AR"]),
        (M + "test_qwen3", []),
        (M + "test_qwen3", ["--unfused"]),
    ]
    mega = [
        ("triton_dist.mega_triton_kernel.test.ops.test_ops", []),
        (M + "test_qwen3", []),
        (M + "test_qwen3", ["--unfused"]),
    ]
    layers = [
        (T + "test_tp_mlp", ["--check", "--verify-iters", 2, "--stress"]),
        (T + "test_tp_attn", ["--check", "--verify-iters", 1]),
        (T + "test_moe_utils", ["--check", "--verify-iters", 2]),
        (T + "test_ag_moe", ["--check", "--verify-iters", 2, "--stress"]),
        (T + "test_moe_reduce_rs", ["--check", "--verify-iters", 2, "--stress"]),
        (T + "test_ep_moe_inference", ["--check", "--verify-iters", 2]),
    ] + ([(T + "test_ep_moe_inference", ["--check", "--verify-iters", 1, "--fp8"])] if torch.cuda.is_available() else []) + [
        (T + "test_decode_attn", ["--check", "--verify-iters", 2]),
        (T + "test_e2e_inference", ["--gen-len", 6]),
    ]
    prims = [
        (T + "test_ring_put", ["--stress"]),
        (T + "test_notify", []),
        (T + "test_distributed_wait", []),
        (T + "test_common_ops", []),
        (T + "test_multimem_ld_reduce", []),
        (T + "test_nvshmem_api", []),
        (T + "test_utils", []),
        (T + "test_all_gather", ["--check", "--verify-iters", 2, "--stress"]),
        (T + "test_ag_small_msg", ["--check", "--verify-iters", 2]),
        (T + "test_pp", ["--stress"]),
    ]
    return {"core": core, "moe": moe, "models": models, "mega": mega, "layers": layers, "prims": prims,
            "all": core + moe + models + mega[:1] + layers + prims}[suite]


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--suite", default="all", choices=["core", "moe", "models", "mega", "layers", "prims", "all"])
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    ok = run_suite(cases(args.suite, world))
    sys.exit(0 if ok else 1)
