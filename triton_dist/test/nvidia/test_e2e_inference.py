"""Engine.serve end to end: prefill + CUDA-graph decode of a small random-init Qwen3 in every backend; the generated
token ids must agree with the torch backend (mirrors reference test/nvidia/test_e2e_inference.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_e2e_inference --model Qwen/Qwen3-32B --bsz 8 --gen-len 64"""
import argparse

import torch

from triton_dist.models import DenseLLM, Engine, ModelConfig
from triton_dist.test.common import all_ok, finish, setup
from triton_dist.test.nvidia.test_tp_e2e import build_hf
from triton_dist.utils import dist_print

if __name__ == "__main__":
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--model", default="", help="HF model name (random init from its config when weights are not cached); "
                    "default: a tiny in-memory Qwen3")
    ap.add_argument("--bsz", type=int, default=0)
    ap.add_argument("--prompt-len", type=int, default=16)
    ap.add_argument("--gen-len", type=int, default=8)
    ap.add_argument("--backends", default="torch,triton_dist_AR,triton_dist_gemm_ar,triton_dist")
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    B = args.bsz or world
    if args.model:
        cfg = ModelConfig(model_name=args.model, max_length=args.prompt_len + args.gen_len + 8, dtype=dtype, rank=rank, world_size=world)
        model = DenseLLM(cfg, group)
        vocab = model.vocab_size
    else:
        hf = build_hf("qwen3", dtype, heads=max(4, world), kv_heads=max(2, world))
        cfg = ModelConfig(model_name="hf-in-memory", max_length=args.prompt_len + args.gen_len + 8, dtype=dtype, rank=rank, world_size=world)
        model = DenseLLM(cfg, group, hf_model=hf)
        vocab = 1024
    torch.manual_seed(1)
    prompt = torch.randint(0, vocab, (B, args.prompt_len))
    outs, ok = {}, True
    for backend in args.backends.split(","):
        eng = Engine(cfg, temperature=0.0, group=group, model=model)
        eng.backend = backend
        out = eng.serve(prompt, args.gen_len)
        if backend == "triton_dist" and world > 1:  # batch-sharded output: gather the shards
            import torch.distributed as dist

            parts = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather_object(parts, out)
            out = torch.cat(parts, 0)
        outs[backend] = out
        dist_print(f"[e2e] backend {backend:22s}: {eng.last_decode_ms_per_token:.3f} ms/token (host clock), first ids {out[0, :6].tolist()}")
    ref = outs.get("torch")
    if ref is not None:
        for k, v in outs.items():
            # greedy decoding of a random-init model: bf16 rounding can flip near-ties, so require the first step to agree
            # everywhere and most of the sequence to agree
            agree = float((v == ref).float().mean())
            good = bool(torch.equal(v[:, 0], ref[:, 0])) and agree > 0.7
            if not good:
                dist_print(f"[e2e] backend {k}: token agreement with torch {agree:.2f}", allowed_ranks="all", prefix=True)
            ok = ok and good
    ok = all_ok(ok)
    finish(ok)
