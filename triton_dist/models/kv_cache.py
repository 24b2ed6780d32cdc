"""Dense per-layer KV cache (reference ``models/kv_cache.py:29-66``): [L, B, max_len, Hkv/W, D] for K and V.
The attention layer's fused norm+RoPE kernel appends K/V in place at the token positions, so
``update_kv_cache`` only hands out the layer's views (same contract as the reference)."""
from __future__ import annotations

import torch


class KV_Cache:

    def __init__(self, num_layers: int = 32, batch_size: int = 1, max_length: int = 32 * 1024, kv_heads: int = 8,
                 head_dim: int = 128, dtype=torch.bfloat16, world_size: int = 8, device=None) -> None:
        self.num_layers, self.batch_size, self.max_length = num_layers, batch_size, max_length
        self.kv_heads, self.head_dim, self.dtype, self.world_size = kv_heads, head_dim, dtype, world_size
        dev = device or ("cuda" if torch.cuda.is_available() else "cpu")
        local_heads = max(kv_heads // world_size, 1)
        shape = (num_layers, batch_size, max_length, local_heads, head_dim)
        self.k_cache = torch.zeros(shape, device=dev, dtype=dtype)
        self.v_cache = torch.zeros(shape, device=dev, dtype=dtype)
        self.kv_offset = torch.zeros(batch_size, dtype=torch.int32, device=dev)

    def update_kv_cache(self, new_k_cache, new_v_cache, layer_idx: int):
        return self.k_cache[layer_idx], self.v_cache[layer_idx], self.kv_offset

    def rand_fill_kv_cache(self, offset: int):
        shape = self.k_cache[:, :, :offset].shape
        self.k_cache[:, :, :offset].copy_(torch.rand(shape, device=self.k_cache.device, dtype=self.dtype) / 10)
        self.v_cache[:, :, :offset].copy_(torch.rand(shape, device=self.k_cache.device, dtype=self.dtype) / 10)

    def inc_offset(self):
        self.kv_offset += 1

    def clear(self):
        self.kv_offset.zero_()

    def get_kv_len(self):
        return self.kv_offset
