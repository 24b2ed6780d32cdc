"""DenseLLM: tensor-parallel decoder-only transformer (Qwen3 / Llama-3 / Seed-OSS) for inference
(reference ``models/dense.py:53-241``).

Weights come from an HF-format checkpoint (local directory or an in-memory HF module) sharded per layer, or are
random-initialised directly on each GPU (``RANDOM_PARAMS=1`` or no checkpoint available) from the built-in
architecture table. Forward modes (``set_fwd``): torch | triton_dist | triton_dist_AR | triton_dist_gemm_ar.
In the three native modes every kernel on the step is a native sm_100a kernel: embedding gather, RMSNorm with the
residual add fused in, GEMV/tcgen05 linears, fused norm+RoPE+KV-append, split-KV flash decode, fused GEMM+AR /
AG+GEMM / GEMM+RS."""
from __future__ import annotations

import gc
import os
from typing import List, Optional

import torch
import torch.nn.functional as F

from ..kernels.allreduce import AllReduceMethod
from ..kernels.nvidia.elementwise import embedding, linear, make_cos_sin_cache, rmsnorm
from ..layers.nvidia.tp_attn import TP_Attn, _set_cos_sin_cache, layer_norm
from ..layers.nvidia.tp_mlp import TP_MLP, _dev
from .config import ModelConfig, load_arch
from .kv_cache import KV_Cache
from .utils import init_model_cpu, logger, random_params_requested


class DenseLLMLayer:

    def __init__(self, layer_idx: int, group) -> None:
        self.attn: Optional[TP_Attn] = None
        self.mlp = None
        self.input_norm_eps = self.post_norm_eps = 1e-6
        self.input_norm_w = self.post_norm_w = None
        self.layer_idx, self.group = layer_idx, group
        self.native = False

    def _make_mlp(self, rank, world_size):
        return TP_MLP(rank=rank, world_size=world_size, group=self.group)

    def init_parameters(self, hf_layer, rank: int, world_size: int, head_dim: int = 128):
        self.mlp = self._make_mlp(rank, world_size)
        self.mlp._init_parameters(hf_layer.mlp)
        self.attn = TP_Attn(rank=rank, world_size=world_size, group=self.group, head_dim=head_dim)
        self.attn._init_parameters(hf_layer.self_attn)
        self.input_norm_eps = hf_layer.input_layernorm.variance_epsilon
        self.input_norm_w = hf_layer.input_layernorm.weight.detach().to(_dev())
        self.post_norm_eps = hf_layer.post_attention_layernorm.variance_epsilon
        self.post_norm_w = hf_layer.post_attention_layernorm.weight.detach().to(_dev())

    def init_random(self, arch, rank: int, world_size: int, dtype, gen):
        self.mlp = self._make_mlp(rank, world_size)
        self._init_random_mlp(arch, dtype, gen)
        self.attn = TP_Attn(rank=rank, world_size=world_size, group=self.group, head_dim=arch.head_dim)
        self.attn._init_random(arch.hidden_size, arch.num_attention_heads, arch.num_key_value_heads, dtype,
                               qk_norm=arch.qk_norm, qkv_bias=arch.attention_bias, gen=gen)
        self.attn.q_norm_eps = self.attn.k_norm_eps = arch.rms_norm_eps
        self.input_norm_eps = self.post_norm_eps = arch.rms_norm_eps
        self.input_norm_w = torch.ones(arch.hidden_size, device=_dev(), dtype=dtype)
        self.post_norm_w = torch.ones(arch.hidden_size, device=_dev(), dtype=dtype)

    def _init_random_mlp(self, arch, dtype, gen):
        self.mlp._init_random(arch.hidden_size, arch.intermediate_size, dtype, gen)

    def set_fwd(self, mode: str = "torch"):
        table = {"torch": "torch_fwd", "triton_dist": "dist_triton_fwd", "triton_dist_AR": "dist_triton_AR_fwd",
                 "triton_dist_gemm_ar": "dist_triton_gemm_ar_fwd"}
        if mode not in table:
            raise ValueError(f"Unsupported mode: {mode}, choose from {list(table)}")
        self.attn.fwd = getattr(self.attn, table[mode])
        self.mlp.fwd = getattr(self.mlp, table[mode])
        self.native = mode != "torch"

    @torch.inference_mode()
    def fwd(self, hidden_states, position_ids, cos_sin_cache, kv_cache: KV_Cache, residual=None):
        """Pre-norm block. Returns (block output before the residual add, running residual): the add is fused into the
        next RMSNorm kernel, so the caller carries (delta, residual) between layers."""
        if residual is None:
            residual = hidden_states
            h = layer_norm(hidden_states, self.input_norm_eps, self.input_norm_w)
        else:
            h = rmsnorm(hidden_states, self.input_norm_w, self.input_norm_eps, residual=residual)
        h = self.attn.fwd(h, position_ids, cos_sin_cache, kv_cache, self.layer_idx)
        h = rmsnorm(h, self.post_norm_w, self.post_norm_eps, residual=residual)
        h = self.mlp.fwd(h)
        return h, residual


class DenseLLM:

    layer_cls = DenseLLMLayer
    model_type = "dense"

    def __init__(self, model_config: ModelConfig, group=None, hf_model=None) -> None:
        self.dtype = model_config.dtype
        self.model_name = model_config.model_name
        self.max_length = model_config.max_length
        self.rank, self.world_size, self.group = model_config.rank, model_config.world_size, group
        self.arch = load_arch(self.model_name, model_config.local_only) if hf_model is None else self._arch_from_hf(hf_model)
        a = self.arch
        self.config = a
        self.hidden_size, self.num_heads, self.head_dim = a.hidden_size, a.num_attention_heads, a.head_dim
        self.num_key_value_heads = a.num_key_value_heads
        self.num_key_value_groups = self.num_heads // self.num_key_value_heads
        self.max_position_embeddings, self.rope_theta = a.max_position_embeddings, a.rope_theta
        self.layers: List[DenseLLMLayer] = []
        self.init_parameters(hf_model)
        self.set_fwd()
        self.use_ar = False

    @staticmethod
    def _arch_from_hf(hf_model):
        from types import SimpleNamespace

        from .config import _cfg

        c = hf_model.config
        d = _cfg()
        for k in d:
            if hasattr(c, k) and getattr(c, k) is not None:
                d[k] = getattr(c, k)
        for k in ("hidden_size", "num_hidden_layers", "num_attention_heads", "num_key_value_heads", "intermediate_size"):
            d[k] = getattr(c, k)
        if getattr(c, "head_dim", None) is None:
            d["head_dim"] = c.hidden_size // c.num_attention_heads
        rp = getattr(c, "rope_parameters", None)
        if isinstance(rp, dict) and rp.get("rope_theta"):
            d["rope_theta"] = rp["rope_theta"]
        d["qk_norm"] = hasattr(hf_model.model.layers[0].self_attn, "q_norm")
        d["model_type"] = c.model_type
        return SimpleNamespace(**d)

    def set_fwd(self, mode: str = "torch"):
        self.mode = mode
        for layer in self.layers:
            layer.set_fwd(mode)

    # ------------------------------------------------------------------ parameters
    def init_parameters(self, hf_model=None):
        a = self.arch
        use_random = hf_model is None and (random_params_requested() or not os.path.isdir(self.model_name))
        if use_random:
            logger.log(f"random-init {self.model_name} shard {self.rank}/{self.world_size} on the device", "info")
            gen = torch.Generator(device=_dev()).manual_seed(20240607)  # same seed on all ranks: replicated tensors agree
            self.embed_tokens = (torch.randn(a.vocab_size, a.hidden_size, device=_dev(), generator=gen) * 0.02).to(self.dtype)
            self.lm_head = self.embed_tokens if a.tie_word_embeddings else \
                (torch.randn(a.vocab_size, a.hidden_size, device=_dev(), generator=gen) * 0.02).to(self.dtype)
            self.norm_weight = torch.ones(a.hidden_size, device=_dev(), dtype=self.dtype)
            self.norm_variance_epsilon = a.rms_norm_eps
            self.cos_sin_cache = make_cos_sin_cache(a.head_dim, self.max_length, a.rope_theta, device=_dev())
            shard_gen = torch.Generator(device=_dev()).manual_seed(1000 + self.rank)
            for idx in range(a.num_hidden_layers):
                layer = self.layer_cls(idx, self.group)
                layer.init_random(a, self.rank, self.world_size, self.dtype, shard_gen)
                self.layers.append(layer)
        else:
            if hf_model is None:
                hf_model = init_model_cpu(self.model_name, dtype=self.dtype)
            self.embed_tokens = hf_model.model.embed_tokens.weight.detach().to(_dev())
            self.lm_head = hf_model.lm_head.weight.detach().to(_dev())
            self.norm_weight = hf_model.model.norm.weight.detach().to(_dev())
            self.norm_variance_epsilon = hf_model.model.norm.variance_epsilon
            inv_freq = hf_model.model.rotary_emb.inv_freq.to(_dev())
            assert inv_freq.numel() * 2 == a.head_dim
            self.cos_sin_cache = _set_cos_sin_cache(inv_freq, max_length=self.max_length)
            for idx, hf_layer in enumerate(hf_model.model.layers):
                layer = self.layer_cls(idx, self.group)
                layer.init_parameters(hf_layer=hf_layer, rank=self.rank, world_size=self.world_size, head_dim=a.head_dim)
                self.layers.append(layer)
            gc.collect()
        self.num_layers = len(self.layers)

    # ------------------------------------------------------------------ distributed contexts (one per model, shared by layers)
    def init_triton_dist_ctx(self, max_M: int = 4096):
        l0 = self.layers[0]
        l0.attn._init_ctx(max_M=max_M)
        l0.mlp._init_ctx(max_M=max_M)
        for layer in self.layers[1:]:
            layer.attn.ag_ctx, layer.attn.rs_ctx = l0.attn.ag_ctx, l0.attn.rs_ctx
            layer.mlp.ag_ctx, layer.mlp.rs_ctx = l0.mlp.ag_ctx, l0.mlp.rs_ctx
        self.use_ar = False

    def init_triton_dist_AR_ctx(self, max_M: int = 128, ar_method: AllReduceMethod = AllReduceMethod.Unknown):
        l0 = self.layers[0]
        l0.attn._init_AR_ctx(max_M=max_M, method=ar_method, dtype=self.dtype)
        l0.mlp._init_AR_ctx(max_M=max_M, method=ar_method, dtype=self.dtype)
        for layer in self.layers[1:]:
            layer.attn.ar_ctx, layer.attn.ar_method = l0.attn.ar_ctx, l0.attn.ar_method
            layer.mlp.ar_ctx, layer.mlp.ar_method = l0.mlp.ar_ctx, l0.mlp.ar_method
        self.use_ar = True

    def init_triton_dist_gemm_ar_ctx(self, max_M: int = 4096):
        l0 = self.layers[0]
        l0.attn._init_gemm_ar_ctx(max_M=max_M, dtype=self.dtype)
        l0.mlp._init_gemm_ar_ctx(max_M=max_M, dtype=self.dtype)
        for layer in self.layers[1:]:
            layer.attn.gemm_ar_ctx = l0.attn.gemm_ar_ctx
            layer.mlp.gemm_ar_ctx = l0.mlp.gemm_ar_ctx
        self.use_ar = True

    def finalize(self):
        if self.layers:
            self.layers[0].attn.finalize()
            self.layers[0].mlp.finalize()

    # ------------------------------------------------------------------ forward
    @torch.inference_mode()
    def inference(self, input_ids: torch.LongTensor, position_ids: torch.LongTensor, kv_cache: KV_Cache, wo_lm_head: bool = False):
        bsz, seq_len = input_ids.shape
        hidden = embedding(input_ids, self.embed_tokens)
        residual = None
        for layer in self.layers:
            hidden, residual = layer.fwd(hidden, position_ids, self.cos_sin_cache, kv_cache, residual)
        hidden = rmsnorm(hidden, self.norm_weight, self.norm_variance_epsilon, residual=residual)
        if seq_len > 1:
            hidden = hidden[:, -1:].contiguous()
        if wo_lm_head:
            return hidden
        if self.mode == "torch":
            return F.linear(hidden, self.lm_head).float()
        return linear(hidden, self.lm_head).float()
