"""Inter-node flavour of the SP all-gather attention (reference ``kernels/nvidia/sp_ag_attention_inter_node.py``).
This framework targets one NVSwitch node, where "inter-node" ranks do not exist; the entry points are kept so that code
written against the reference imports and runs on a single node (they forward to the intra-node implementation) and
fail clearly when asked for more than one node."""
from __future__ import annotations

from .sp_ag_attention_intra_node import (SPAllGatherAttentionContextIntraNode, create_sp_ag_attention_context_intra_node,
                                         fused_sp_ag_attn_intra_node)

SPAllGatherAttentionContextInterNode = SPAllGatherAttentionContextIntraNode


def create_sp_ag_attention_context_inter_node(batch_size, q_head, kv_head, max_seqlen_k, max_q_shard_len, head_dim, input_dtype,
                                              output_dtype, rank, local_rank, world_size, local_world_size, device=None):
    if world_size != local_world_size:
        raise NotImplementedError("single NVSwitch node only: world_size must equal local_world_size")
    return create_sp_ag_attention_context_intra_node(batch_size, q_head, kv_head, max_seqlen_k, max_q_shard_len, head_dim,
                                                     input_dtype, output_dtype, rank, world_size, device)


def fused_sp_ag_attn_inter_node(ctx, q_shard, k_shard, v_shard, output, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                                rank, world_size, local_world_size=None, is_causal=True, enable_zig_zag=True):
    if local_world_size is not None and local_world_size != world_size:
        raise NotImplementedError("single NVSwitch node only")
    return fused_sp_ag_attn_intra_node(ctx, q_shard, k_shard, v_shard, output, cu_seqlens_q, cu_seqlens_k, max_seqlen_q,
                                       max_seqlen_k, rank, world_size, is_causal, enable_zig_zag)
