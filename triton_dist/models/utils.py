"""Model helpers: logger, sampling, weight initialisation (reference ``models/utils.py:45-128``)."""
from __future__ import annotations

import os
import random
import sys

import numpy as np
import torch

from ..kernels.nvidia.elementwise import argmax


class MyLogger:
    """Levelled stdout logger; ``TRITON_DIST_DEBUG=1`` enables debug lines (reference :45-70)."""

    COLORS = {"info": "\033[0m", "success": "\033[92m", "warning": "\033[93m", "error": "\033[91m", "debug": "\033[94m"}

    def __init__(self):
        self.debug_enabled = os.environ.get("TRITON_DIST_DEBUG", "0").lower() in ("1", "true", "yes")
        self.rank = int(os.environ.get("RANK", "0"))

    def log(self, msg: str, level: str = "info"):
        if level == "debug" and not self.debug_enabled:
            return
        if self.rank != 0 and level not in ("error",):
            return
        color = self.COLORS.get(level, "") if sys.stdout.isatty() else ""
        end = "\033[0m" if color else ""
        print(f"{color}[{level.upper()}] {msg}{end}", flush=True)


logger = MyLogger()


def seed_everything(seed: int):
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    np.random.seed(seed)
    random.seed(seed)


@torch.inference_mode()
def sample_token(logits: torch.Tensor, temperature: float = 0.6, top_p: float = 0.95, top_k: int = -1) -> torch.Tensor:
    """Greedy (native argmax kernel) when temperature == 0, else nucleus sampling. Returns [B, 1]."""
    if temperature == 0.0:
        return argmax(logits).unsqueeze(-1)
    assert top_k == -1, "top_k sampling is not supported"
    probs = (logits.float() / temperature).softmax(dim=-1)
    sp, si = probs.sort(dim=-1, descending=True)
    keep = sp.cumsum(dim=-1) - sp < top_p
    sp = torch.where(keep, sp, torch.zeros_like(sp))
    choice = torch.multinomial(sp / sp.sum(dim=-1, keepdim=True), 1)
    return si.gather(-1, choice)


def random_params_requested() -> bool:
    return os.environ.get("RANDOM_PARAMS", "0").lower() in ("1", "true", "yes")


@torch.no_grad()
def init_model_cpu(model_name: str, dtype: torch.dtype):
    """Load an HF checkpoint (local directory or HF cache) on the host; sharding to the GPU happens per layer."""
    from transformers import AutoModelForCausalLM

    return AutoModelForCausalLM.from_pretrained(model_name, dtype=dtype, local_files_only=True)
