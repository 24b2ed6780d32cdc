"""Layers (reference ``layers/nvidia/__init__.py:25-30``)."""
from .ep_a2a_layer import EPAll2AllLayer
from .gemm_allreduce_layer import GemmARLayer
from .low_latency_allgather_layer import AllGatherLayer
from .p2p import CommOp
from .sp_flash_decode_layer import SpGQAFlashDecodeAttention
from .tp_attn import TP_Attn
from .tp_mlp import TP_MLP
from .tp_moe import TP_MoE

__all__ = ["AllGatherLayer", "CommOp", "EPAll2AllLayer", "GemmARLayer", "SpGQAFlashDecodeAttention", "TP_Attn", "TP_MLP", "TP_MoE"]
