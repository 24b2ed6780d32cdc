"""SpGQAFlashDecodeAttention: decode attention over a KV cache sharded across ranks on the sequence
dimension (reference ``layers/nvidia/sp_flash_decode_layer.py:44-185``).

Per step: (1) split-KV flash decode over this rank's KV shard keeping the log-sum-exp
(``out | lse``, fp32), (2) low-latency all-gather of the W partials (LL packets, NVLS multimem when
bound), (3) log-sum-exp merge across ranks. Three launches, no host synchronisation."""
from __future__ import annotations

import torch

from ...kernels.nvidia.flash_decode import gqa_fwd_batch_decode_intra_rank, inter_rank_combine
from ...utils import symm_tensor
from .low_latency_allgather_layer import AllGatherLayer


class SpGQAFlashDecodeAttention(torch.nn.Module):

    def __init__(self, rank, node, num_ranks, num_nodes, num_q_heads, num_kv_heads, q_head_dim, v_head_dim, page_size=1,
                 scale=1, soft_cap=0, max_allowed_batch=1, thrink_buffer_threshold=500, stages=20):
        super().__init__()
        assert soft_cap == 0, "soft_cap is not supported"
        self.rank, self.node, self.num_ranks, self.num_nodes = rank, node, num_ranks, num_nodes
        self.num_q_heads, self.num_kv_heads = num_q_heads, num_kv_heads
        self.q_head_dim, self.v_head_dim = q_head_dim, v_head_dim
        self.page_size, self.scale, self.soft_cap = page_size, scale, soft_cap
        self.kv_split = 32
        self.max_allowed_batch = max_allowed_batch
        self.workspace = None
        per_rank = max_allowed_batch * num_q_heads * (v_head_dim + 1) * 4
        self.ag_layer = AllGatherLayer(num_nodes, num_ranks, rank, max_buffer_size=4 * per_rank * num_ranks)
        self.ag_buffer = symm_tensor((num_ranks, max_allowed_batch, num_q_heads, v_head_dim + 1), torch.float32)

    def finalize(self):
        self.ag_layer.finalize()

    def forward(self, q, k_cache, v_cache, global_kv_lens, block_table):
        """q [B, Hq, D] (replicated); k_cache/v_cache: this rank's shard [pages, page_size, Hkv, D];
        global_kv_lens [W, B] int32: shard length per rank; block_table [B, max_pages] for the local shard."""
        batch = q.shape[0]
        assert global_kv_lens.shape == (self.num_ranks, batch) and batch <= self.max_allowed_batch
        buf = self.ag_buffer[:, :batch] if batch == self.max_allowed_batch else None
        local = gqa_fwd_batch_decode_intra_rank(q, k_cache, v_cache, self.workspace, [1] * batch, global_kv_lens[self.rank],
                                                block_table, self.scale, kv_split=self.kv_split,
                                                output_combine=self.ag_buffer[self.rank, :batch] if buf is not None else None)
        if buf is None:  # smaller batch than the buffer was sized for: gather a compact staging view
            stage = self.ag_buffer.view(-1)[:self.num_ranks * local.numel()].view(self.num_ranks, *local.shape)
            stage[self.rank].copy_(local)
            buf = stage
        if self.num_ranks > 1:
            mode = self.ag_layer.forward_push_2d_ll_multimem if q.is_cuda else self.ag_layer.forward_pull
            mode(buf)
        return inter_rank_combine(buf, dtype=q.dtype)
