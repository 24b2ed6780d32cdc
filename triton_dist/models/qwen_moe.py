"""Qwen3MoE: tensor-parallel Qwen3 mixture-of-experts model (reference ``models/qwen_moe.py``).

Attention is the dense TP_Attn; the MLP of every layer is a TP_MoE (experts replicated across ranks, each expert's
FFN dimension sharded): AG + grouped GEMM -> SiLU*up -> grouped GEMM + top-k reduce + reduce-scatter."""
from __future__ import annotations

import torch

from ..layers.nvidia.tp_moe import TP_MoE
from .dense import DenseLLM, DenseLLMLayer


class Qwen3MoELayer(DenseLLMLayer):

    def _make_mlp(self, rank, world_size):
        return TP_MoE(rank=rank, world_size=world_size, group=self.group)

    def _init_random_mlp(self, arch, dtype, gen):
        self.mlp._init_random(arch.hidden_size, arch.moe_intermediate_size, arch.num_experts, arch.num_experts_per_tok, dtype,
                              gen, norm_topk_prob=arch.norm_topk_prob)


class Qwen3MoE(DenseLLM):
    layer_cls = Qwen3MoELayer
    model_type = "moe"

    def init_triton_dist_ctx(self, max_M: int = 4096):
        l0 = self.layers[0]
        l0.attn._init_ctx(max_M=max_M)
        l0.mlp._init_ctx(max_M=max_M)
        for layer in self.layers[1:]:
            layer.attn.ag_ctx, layer.attn.rs_ctx = l0.attn.ag_ctx, l0.attn.rs_ctx
            layer.mlp.share_ctx(l0.mlp)
        self.use_ar = False

    def init_triton_dist_AR_ctx(self, max_M: int = 128, ar_method=None):
        from ..kernels.allreduce import AllReduceMethod

        method = ar_method if ar_method is not None else AllReduceMethod.Unknown
        l0 = self.layers[0]
        l0.attn._init_AR_ctx(max_M=max_M, method=method, dtype=self.dtype)
        l0.mlp._init_AR_ctx(max_M=max_M, method=method, dtype=self.dtype)
        for layer in self.layers[1:]:
            layer.attn.ar_ctx, layer.attn.ar_method = l0.attn.ar_ctx, l0.attn.ar_method
            layer.mlp.share_ctx(l0.mlp)
        self.use_ar = True

    def init_triton_dist_gemm_ar_ctx(self, max_M: int = 4096):
        l0 = self.layers[0]
        l0.attn._init_gemm_ar_ctx(max_M=max_M, dtype=self.dtype)
        l0.mlp._init_gemm_ar_ctx(max_M=max_M, dtype=self.dtype)
        for layer in self.layers[1:]:
            layer.attn.gemm_ar_ctx = l0.attn.gemm_ar_ctx
            layer.mlp.share_ctx(l0.mlp)
        self.use_ar = True
