"""Model configuration (reference ``models/config.py:30-37``) plus the architecture table of the checkpoints
the reference supports (``models/__init__.py:33-42``). The table makes random-init runs possible without network
access; a local HF checkpoint directory (``model_name`` = path) is read through ``transformers.AutoConfig``."""
from __future__ import annotations

import dataclasses
import os
from types import SimpleNamespace

import torch


@dataclasses.dataclass
class ModelConfig:
    model_name: str = "Qwen/Qwen3-32B"
    max_length: int = 4096
    dtype: torch.dtype = torch.bfloat16
    local_only: bool = False
    rank: int = 0
    world_size: int = 8


def _cfg(**kw):
    base = dict(head_dim=128, rms_norm_eps=1e-6, rope_theta=1e6, vocab_size=151936, tie_word_embeddings=False,
                attention_bias=False, qk_norm=True, max_position_embeddings=40960, num_experts=0, num_experts_per_tok=0,
                moe_intermediate_size=0, norm_topk_prob=True, model_type="qwen3")
    base.update(kw)
    return base


KNOWN_ARCHS = {
    "Qwen/Qwen3-0.6B": _cfg(hidden_size=1024, num_hidden_layers=28, num_attention_heads=16, num_key_value_heads=8,
                             intermediate_size=3072, tie_word_embeddings=True),
    "Qwen/Qwen3-8B": _cfg(hidden_size=4096, num_hidden_layers=36, num_attention_heads=32, num_key_value_heads=8,
                           intermediate_size=12288),
    "Qwen/Qwen3-14B": _cfg(hidden_size=5120, num_hidden_layers=40, num_attention_heads=40, num_key_value_heads=8,
                            intermediate_size=17408),
    "Qwen/Qwen3-32B": _cfg(hidden_size=5120, num_hidden_layers=64, num_attention_heads=64, num_key_value_heads=8,
                            intermediate_size=25600),
    "Qwen/Qwen3-30B-A3B": _cfg(hidden_size=2048, num_hidden_layers=48, num_attention_heads=32, num_key_value_heads=4,
                                intermediate_size=6144, num_experts=128, num_experts_per_tok=8, moe_intermediate_size=768,
                                model_type="qwen3_moe"),
    "Qwen/Qwen3-235B-A22B": _cfg(hidden_size=4096, num_hidden_layers=94, num_attention_heads=64, num_key_value_heads=4,
                                  intermediate_size=12288, num_experts=128, num_experts_per_tok=8, moe_intermediate_size=1536,
                                  model_type="qwen3_moe"),
    "meta-llama/Meta-Llama-3-70B": _cfg(hidden_size=8192, num_hidden_layers=80, num_attention_heads=64, num_key_value_heads=8,
                                        intermediate_size=28672, vocab_size=128256, rope_theta=5e5, rms_norm_eps=1e-5,
                                        qk_norm=False, max_position_embeddings=8192, model_type="llama"),
    "meta-llama/Meta-Llama-3-8B": _cfg(hidden_size=4096, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=8,
                                       intermediate_size=14336, vocab_size=128256, rope_theta=5e5, rms_norm_eps=1e-5,
                                       qk_norm=False, max_position_embeddings=8192, model_type="llama"),
    "ByteDance-Seed/Seed-OSS-36B-Instruct": _cfg(hidden_size=5120, num_hidden_layers=64, num_attention_heads=80,
                                                 num_key_value_heads=8, intermediate_size=27648, vocab_size=155136,
                                                 rope_theta=1e7, attention_bias=True, qk_norm=False,
                                                 max_position_embeddings=524288, model_type="seed_oss"),
}


def load_arch(model_name: str, local_only: bool = False) -> SimpleNamespace:
    """Architecture hyper-parameters for ``model_name`` (table entry, or a local HF checkpoint directory)."""
    if model_name in KNOWN_ARCHS:
        return SimpleNamespace(**KNOWN_ARCHS[model_name])
    if os.path.isdir(model_name):
        from transformers import AutoConfig

        c = AutoConfig.from_pretrained(model_name, local_files_only=True)
        d = _cfg()
        for k in d:
            if hasattr(c, k):
                d[k] = getattr(c, k)
        for k in ("hidden_size", "num_hidden_layers", "num_attention_heads", "num_key_value_heads", "intermediate_size"):
            d[k] = getattr(c, k)
        if getattr(c, "head_dim", None) is None:
            d["head_dim"] = c.hidden_size // c.num_attention_heads
        rp = getattr(c, "rope_parameters", None)
        if isinstance(rp, dict) and "rope_theta" in rp:
            d["rope_theta"] = rp["rope_theta"]
        d["qk_norm"] = c.model_type in ("qwen3", "qwen3_moe")
        d["model_type"] = c.model_type
        return SimpleNamespace(**d)
    raise ValueError(f"unknown model {model_name!r}: not in the built-in table {list(KNOWN_ARCHS)} and not a local directory")
