"""End-to-end TP model check against HuggingFace (mirrors reference test/nvidia/test_tp_e2e.py --check):
a small random Qwen3 (or Llama / Qwen3-MoE) is built identically on every rank, sharded into DenseLLM / Qwen3MoE, and
prefill + decode logits of every backend are compared with the HF model's.
    torchrun --nproc-per-node 2 -m triton_dist.test.nvidia.test_tp_e2e --model qwen3 --check"""
import argparse

import torch
import torch.distributed as dist

from triton_dist.models import DenseLLM, KV_Cache, ModelConfig, Qwen3MoE
from triton_dist.test.common import all_ok, finish, logits_close, setup
from triton_dist.utils import dist_print


def build_hf(kind: str, dtype, heads: int, kv_heads: int):
    import transformers as tf

    torch.manual_seed(1234)
    common = dict(hidden_size=512, num_hidden_layers=2, num_attention_heads=heads, num_key_value_heads=kv_heads, head_dim=128,
                  vocab_size=1024, max_position_embeddings=512, rms_norm_eps=1e-6, tie_word_embeddings=False)
    if kind == "qwen3":
        cfg = tf.Qwen3Config(intermediate_size=1024, **common)
        model = tf.Qwen3ForCausalLM(cfg)
    elif kind == "llama":
        cfg = tf.LlamaConfig(intermediate_size=1024, **common)
        model = tf.LlamaForCausalLM(cfg)
    elif kind == "qwen3_moe":
        cfg = tf.Qwen3MoeConfig(intermediate_size=1024, moe_intermediate_size=256, num_experts=8, num_experts_per_tok=2,
                                norm_topk_prob=True, decoder_sparse_step=1, mlp_only_layers=[], **common)
        model = tf.Qwen3MoeForCausalLM(cfg)
    else:
        raise ValueError(kind)
    return model.to(dtype).eval()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3", choices=["qwen3", "llama", "qwen3_moe"])
    ap.add_argument("--bsz", type=int, default=2)
    ap.add_argument("--seq", type=int, default=24)
    ap.add_argument("--decode-steps", type=int, default=3)
    ap.add_argument("--modes", default="torch,triton_dist_AR,triton_dist_gemm_ar,triton_dist")
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--corrupt", action="store_true",
                    help="self-test of the comparison: 1 %% of the LM-head rows are swapped, the run passes only if every mode is "
                         "reported as a MISMATCH")
    args = ap.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    heads = max(4, world)
    hf = build_hf(args.model, dtype, heads=heads, kv_heads=max(world, 2) if args.model != "qwen3_moe" else max(world, 2))
    cfg = ModelConfig(model_name="hf-in-memory", max_length=128, dtype=dtype, rank=rank, world_size=world)
    cls = Qwen3MoE if args.model == "qwen3_moe" else DenseLLM
    model = cls(cfg, group, hf_model=hf)
    hf = hf.to(dev)
    if args.corrupt:  # a small, structured fault: ten vocabulary rows of the LM head exchanged with their neighbours
        model.lm_head = model.lm_head.clone()  # on CPU the model's LM head aliases the HF module's weight
        rows = torch.arange(0, 1000, 100, device=model.lm_head.device)
        tmp = model.lm_head[rows].clone()
        model.lm_head[rows] = model.lm_head[rows + 1]
        model.lm_head[rows + 1] = tmp
    bsz = (args.bsz + world - 1) // world * world
    torch.manual_seed(7)
    ids = torch.randint(0, 1024, (bsz, args.seq), device=dev)
    with torch.inference_mode():
        ref_all = hf(ids).logits.float()  # teacher forcing: logits at every position
    ok = True
    frac = 0.5 if (args.model == "qwen3_moe" and dtype != torch.float32) else None  # see logits_close: expert flips in 16 bit
    for mode in args.modes.split(","):
        if mode == "triton_dist" and args.model == "qwen3_moe" and world == 1 and dev == "cpu":
            pass
        kv = KV_Cache(num_layers=model.num_layers, batch_size=bsz, max_length=128, kv_heads=model.num_key_value_heads,
                      head_dim=model.head_dim, dtype=dtype, world_size=world)
        n_pre = args.seq - args.decode_steps
        # prefill with the torch path (as the engine does), then decode with `mode`
        model.set_fwd("torch")
        pos = torch.arange(n_pre, device=dev)[None, :].expand(bsz, -1)
        logits = model.inference(ids[:, :n_pre], pos, kv)
        good = logits_close(logits[:, -1], ref_all[:, n_pre - 1], dtype, f"mode {mode} prefill", frac)
        kv.kv_offset.fill_(n_pre)
        if mode == "triton_dist":
            model.set_fwd("triton_dist")
            model.init_triton_dist_ctx(max_M=bsz)
        elif mode == "triton_dist_AR":
            model.set_fwd("triton_dist_AR")
            model.init_triton_dist_AR_ctx(max_M=bsz)
        elif mode == "triton_dist_gemm_ar":
            model.set_fwd("triton_dist_gemm_ar")
            model.init_triton_dist_gemm_ar_ctx(max_M=bsz)
        for step in range(args.decode_steps):
            p = n_pre + step
            tok = ids[:, p:p + 1]
            if mode == "triton_dist":
                tok = tok.split(bsz // world, dim=0)[rank]
            pos = torch.full((bsz, 1), p, device=dev, dtype=torch.long)
            lg = model.inference(tok, pos, kv)[:, -1]
            ref = ref_all[:, p]
            if mode == "triton_dist":
                ref = ref.split(bsz // world, dim=0)[rank]
            step_ok = logits_close(lg, ref, dtype, f"mode {mode} decode step {step}", frac)
            good = good and step_ok
            kv.inc_offset()
        model.finalize()
        good = all_ok(good)
        dist_print(f"[{args.model}] mode {mode:22s}: {'ok' if good else 'MISMATCH'}")
        ok = ok and (good if not args.corrupt else not good)
    if args.corrupt:
        dist_print("corrupted LM head detected by every mode" if ok else "a corrupted LM head went UNNOTICED")
    finish(ok)
