"""Decode latency of the megakernel vs the CUDA-graph engine paths (mirrors reference
mega_triton_kernel/test/models/bench_qwen3.py; published numbers docs/mega_triton_kernel.md:29-39).
    torchrun --nproc-per-node 8 -m triton_dist.mega_triton_kernel.test.models.bench_qwen3 --model Qwen/Qwen3-32B --ctx 512
Reports ms/step (device-timed, max over ranks) for: torch eager, torch + CUDA graph, triton_dist_AR + graph,
triton_dist_gemm_ar + graph, and the megakernel; plus the HBM roofline of the step."""
import argparse
import json

import torch
import torch.distributed as dist

from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
from triton_dist.models import DenseLLM, KV_Cache, ModelConfig
from triton_dist.test.common import finish, setup
from triton_dist.utils import dist_print, measured_peaks, perf_func_max_over_ranks


def weight_bytes_per_step(dense: DenseLLM) -> int:
    n = dense.lm_head.numel() * dense.lm_head.element_size()
    for layer in dense.layers:
        for w in (layer.attn.wqkv, layer.attn.wo, layer.mlp.gate_up_proj, layer.mlp.down_proj):
            n += w.numel() * w.element_size()
    return n


def bench_engine_mode(dense: DenseLLM, mode: str, bsz: int, ctx: int, iters: int, graph: bool) -> float:
    kv = KV_Cache(num_layers=dense.num_layers, batch_size=bsz, max_length=dense.max_length, kv_heads=dense.num_key_value_heads,
                  head_dim=dense.head_dim, dtype=dense.dtype, world_size=dense.world_size)
    kv.rand_fill_kv_cache(ctx)
    kv.kv_offset.fill_(ctx)
    dense.set_fwd(mode)
    if mode == "triton_dist_AR":
        dense.init_triton_dist_AR_ctx(max_M=bsz)
    elif mode == "triton_dist_gemm_ar":
        dense.init_triton_dist_gemm_ar_ctx(max_M=bsz)
    ids = torch.full((bsz, 1), 1225, dtype=torch.long, device="cuda")
    pos = torch.full((bsz, 1), ctx, dtype=torch.long, device="cuda")
    fn = lambda: dense.inference(ids, pos, kv)  # noqa: E731
    _, ms = perf_func_max_over_ranks(fn, iters, 3, use_graph=graph)
    if mode != "torch":
        dense.finalize()
    dense.set_fwd("torch")
    return ms


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Qwen/Qwen3-32B")
    ap.add_argument("--bsz", type=int, default=1)
    ap.add_argument("--ctx", type=int, default=512)
    ap.add_argument("--max-length", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--modes", default="torch,torch_graph,triton_dist_AR,triton_dist_gemm_ar,mega")
    ap.add_argument("--json", default="")
    ap.add_argument("--profile", action="store_true", help="megakernel: per-task timeline of one step (summary in the JSON, chrome "
                    "trace in prof/mega_rank<r>.json)")
    args = ap.parse_args()
    group, rank, world = setup()
    cfg = ModelConfig(model_name=args.model, max_length=args.max_length, dtype=torch.bfloat16, rank=rank, world_size=world)
    dense = DenseLLM(cfg, group)
    res = {"model": args.model, "tp": world, "bsz": args.bsz, "ctx": args.ctx}
    modes = args.modes.split(",")
    with torch.inference_mode():
        if "torch" in modes:
            res["torch_eager_ms"] = bench_engine_mode(dense, "torch", args.bsz, args.ctx, args.iters, graph=False)
        if "torch_graph" in modes:
            res["torch_graph_ms"] = bench_engine_mode(dense, "torch", args.bsz, args.ctx, args.iters, graph=True)
        if "triton_dist_AR" in modes:
            res["triton_dist_AR_graph_ms"] = bench_engine_mode(dense, "triton_dist_AR", args.bsz, args.ctx, args.iters, graph=True)
        if "triton_dist_gemm_ar" in modes:
            res["triton_dist_gemm_ar_graph_ms"] = bench_engine_mode(dense, "triton_dist_gemm_ar", args.bsz, args.ctx, args.iters, graph=True)
        if "mega" in modes:
            builder = ModelBuilder(rank, world, world)
            mega = Qwen3Model(args.bsz, cfg, builder, dense=dense)
            mega.set_kv_len(args.ctx, fill_random=True)
            ids = torch.full((args.bsz,), 1225, dtype=torch.long, device="cuda")

            def step():
                mega.kv_cache.set_position(args.ctx)
                return mega.mega_forwrad(ids)

            _, ms = perf_func_max_over_ranks(step, args.iters, 3)
            res["mega_ms"] = ms
            res["mega_graph"] = builder.info()
            if args.profile:
                import os

                builder.enable_profiling(True)
                step()
                torch.cuda.synchronize()
                res["mega_task_profile"] = builder.profile_summary()
                os.makedirs("prof", exist_ok=True)
                builder.dump_trace(f"prof/mega_rank{rank}.json")
                builder.enable_profiling(False)
            mega.finalize()
    wb = weight_bytes_per_step(dense)
    pk = measured_peaks()
    res["weight_bytes_per_step_per_rank"] = wb
    res["hbm_roofline_ms"] = wb / (pk["hbm_gbs"] * 1e6)
    if "mega_ms" in res:
        res["mega_fraction_of_hbm_roofline_" + pk["source"]] = res["hbm_roofline_ms"] / res["mega_ms"]
    dist_print("bench_qwen3 " + json.dumps(res))
    if args.json and rank == 0:
        with open(args.json, "w") as f:
            json.dump(res, f, indent=1)
    finish(True)
