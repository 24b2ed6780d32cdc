"""Persistent decode megakernel ("MegaTritonKernel" in the reference, python/triton_dist/mega_triton_kernel):
a whole decode step as ONE cooperative launch of a hand-written sm_100a interpreter kernel over a static task graph.

    builder = ModelBuilder(rank, world_size, local_world_size)
    builder.make_rms_norm(...); builder.make_qkv_proj(...); ...; builder.compile(); builder.run()
"""
from .core.builder import ModelBuilder
from .models.paged_kv_cache import PagedKVCache
from .models.qwen3 import Qwen3Model

__all__ = ["ModelBuilder", "PagedKVCache", "Qwen3Model"]
