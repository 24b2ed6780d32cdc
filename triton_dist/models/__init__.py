"""Models and the inference engine (reference ``models/__init__.py:25-64``)."""
from .config import KNOWN_ARCHS, ModelConfig
from .dense import DenseLLM
from .engine import Engine
from .kv_cache import KV_Cache
from .qwen_moe import Qwen3MoE


class AutoLLM:
    model_mapping = {
        "Qwen/Qwen3-0.6B": DenseLLM,
        "Qwen/Qwen3-8B": DenseLLM,
        "Qwen/Qwen3-14B": DenseLLM,
        "Qwen/Qwen3-32B": DenseLLM,
        "Qwen/Qwen3-30B-A3B": Qwen3MoE,
        "Qwen/Qwen3-235B-A22B": Qwen3MoE,
        "meta-llama/Meta-Llama-3-70B": DenseLLM,
        "meta-llama/Meta-Llama-3-8B": DenseLLM,
        "ByteDance-Seed/Seed-OSS-36B-Instruct": DenseLLM,
    }

    @staticmethod
    def from_pretrained(config: ModelConfig, group=None):
        cls = AutoLLM.model_mapping.get(config.model_name)
        if cls is None:
            from .config import load_arch

            cls = Qwen3MoE if load_arch(config.model_name).num_experts > 0 else DenseLLM
        return cls(config, group)


class AutoTokenizer:

    @staticmethod
    def from_pretrained(model_config: ModelConfig):
        from transformers import AutoTokenizer as HFTokenizer

        return HFTokenizer.from_pretrained(model_config.model_name, use_fast=True, local_files_only=True)


__all__ = ["AutoLLM", "AutoTokenizer", "DenseLLM", "Engine", "KNOWN_ARCHS", "KV_Cache", "ModelConfig", "Qwen3MoE"]
