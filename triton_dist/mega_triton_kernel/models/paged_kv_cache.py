"""Paged KV cache of the megakernel models (reference ``mega_triton_kernel/models/paged_kv_cache.py:28-51``):
per layer [num_pages, page_size, Hkv/W, D] K and V plus a (random by default, as in the reference) block table."""
from __future__ import annotations

import torch


class PagedKVCache:

    def __init__(self, num_layers: int, batch_size: int, max_length: int, kv_heads: int, head_dim: int, page_size: int = 16,
                 dtype=torch.bfloat16, world_size: int = 1, device=None, random_block_table: bool = True):
        dev = device or ("cuda" if torch.cuda.is_available() else "cpu")
        self.num_layers, self.batch_size, self.max_length, self.page_size = num_layers, batch_size, max_length, page_size
        self.local_kv_heads = max(kv_heads // world_size, 1)
        self.pages_per_seq = (max_length + page_size - 1) // page_size
        num_pages = batch_size * self.pages_per_seq
        shape = (num_layers, num_pages, page_size, self.local_kv_heads, head_dim)
        self.k_cache = torch.zeros(shape, dtype=dtype, device=dev)
        self.v_cache = torch.zeros(shape, dtype=dtype, device=dev)
        order = torch.randperm(num_pages, device=dev) if random_block_table else torch.arange(num_pages, device=dev)
        self.block_table = order.to(torch.int32).view(batch_size, self.pages_per_seq).contiguous()
        # position written by the CURRENT step (kv length - 1); the megakernel reads it from the device
        self.positions = torch.zeros(batch_size, dtype=torch.int32, device=dev)

    def layer(self, idx: int):
        return self.k_cache[idx], self.v_cache[idx]

    def rand_fill(self, length: int):
        for b in range(self.batch_size):
            pages = self.block_table[b, :(length + self.page_size - 1) // self.page_size].long()
            self.k_cache[:, pages] = (torch.rand_like(self.k_cache[:, pages]) - 0.5) / 5
            self.v_cache[:, pages] = (torch.rand_like(self.v_cache[:, pages]) - 0.5) / 5
        self.positions.fill_(length)

    def set_position(self, pos: int):
        self.positions.fill_(pos)

    def inc(self):
        self.positions += 1
