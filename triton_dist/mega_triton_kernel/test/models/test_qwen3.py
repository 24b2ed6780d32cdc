"""Megakernel vs the layer-by-layer model (mirrors reference mega_triton_kernel/test/models/test_qwen3.py).
    torchrun --nproc-per-node 2 -m triton_dist.mega_triton_kernel.test.models.test_qwen3 [--unfused]"""
import argparse

import torch
import torch.distributed as dist

from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
from triton_dist.models import DenseLLM, KV_Cache, ModelConfig
from triton_dist.test.common import logits_close, all_ok, finish, setup
from triton_dist.test.nvidia.test_tp_e2e import build_hf
from triton_dist.utils import dist_print

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--bsz", type=int, default=2)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--ctx", type=int, default=20)
    ap.add_argument("--unfused", action="store_true")
    args = ap.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    hf = build_hf("qwen3", dtype, heads=max(4, world), kv_heads=max(2, world))
    cfg = ModelConfig(model_name="hf-in-memory", max_length=128, dtype=dtype, rank=rank, world_size=world)
    dense = DenseLLM(cfg, group, hf_model=hf)
    hf = hf.to(dev)
    B = args.bsz
    torch.manual_seed(3)
    ids = torch.randint(0, 1024, (B, args.ctx + args.steps), device=dev)
    with torch.inference_mode():
        ref_all = hf(ids).logits.float()
    # fill both caches by running the prompt through the layer-by-layer model, token by token, for the mega model we
    # replay the same tokens through the megakernel itself (it is a decode-only kernel)
    builder = ModelBuilder(rank, world, world)
    mega = Qwen3Model(B, cfg, builder, dense=dense, page_size=16, fused=not args.unfused)
    mega.set_kv_len(0)
    ok = True
    for p in range(args.ctx + args.steps):
        logits = mega.mega_forwrad(ids[:, p]).float()
        if p >= args.ctx:
            good = logits_close(logits, ref_all[:, p], dtype, f"megakernel pos {p}")
            good = good and torch.equal(mega.next_ids, logits.argmax(-1))
            ok = ok and good
    ok = all_ok(ok)
    dist_print(f"megakernel graph: {builder.info()}")
    mega.finalize()
    finish(ok)
