"""Kernel library: context + op pairs with the names exported by the reference's
``python/triton_dist/kernels/nvidia/__init__.py:25-87``; every op launches a
hand-written sm_100a kernel from libtdist.so."""
from .allgather import (AllGatherMethod, cp_engine_producer_all_gather_inter_node, cp_engine_producer_all_gather_intra_node,
                        get_auto_all_gather_method)
from .allgather_gemm import ag_gemm, create_ag_gemm_context, gemm_non_persistent, gemm_persistent
from .gemm_allreduce import (create_gemm_ar_context, create_ll_gemm_ar_context, gemm_allreduce_op,
                             low_latency_gemm_allreduce_op)
from .allgather_group_gemm import ag_group_gemm, create_ag_group_gemm_context
from .flash_decode import (get_triton_combine_kv_algo_info, gqa_fwd_batch_decode, gqa_fwd_batch_decode_aot,
                           gqa_fwd_batch_decode_intra_rank, gqa_fwd_batch_decode_intra_rank_aot, gqa_fwd_batch_decode_persistent,
                           gqa_fwd_batch_decode_persistent_aot, kernel_gqa_fwd_batch_decode_split_kv_persistent,
                           kernel_inter_rank_gqa_fwd_batch_decode_combine_kv)
from .gemm_reduce_scatter import create_gemm_rs_context, gemm_rs
from .low_latency_all_to_all import all_to_all_post_process, create_all_to_all_context, fast_all_to_all
from .moe_reduce_rs import create_moe_rs_context
from .sp_ag_attention_inter_node import create_sp_ag_attention_context_inter_node, fused_sp_ag_attn_inter_node
from .sp_ag_attention_intra_node import create_sp_ag_attention_context_intra_node, fused_sp_ag_attn_intra_node
from .low_latency_allgather import (_forward_pull_kernel, _forward_push_2d_kernel, _forward_push_2d_ll_kernel,
                                    _forward_push_2d_ll_multimem_kernel, _forward_push_3d_kernel,
                                    _forward_push_numa_2d_kernel, _forward_push_numa_2d_ll_kernel,
                                    _forward_push_numa_2d_ll_multinode_kernel, create_fast_allgather_context, fast_allgather)

__all__ = [
    "ag_group_gemm", "all_to_all_post_process", "create_ag_group_gemm_context", "create_all_to_all_context", "create_moe_rs_context",
    "create_sp_ag_attention_context_inter_node", "create_sp_ag_attention_context_intra_node", "fast_all_to_all",
    "fused_sp_ag_attn_inter_node", "fused_sp_ag_attn_intra_node", "get_triton_combine_kv_algo_info", "gqa_fwd_batch_decode",
    "gqa_fwd_batch_decode_aot", "gqa_fwd_batch_decode_intra_rank", "gqa_fwd_batch_decode_intra_rank_aot",
    "gqa_fwd_batch_decode_persistent", "gqa_fwd_batch_decode_persistent_aot", "kernel_gqa_fwd_batch_decode_split_kv_persistent",
    "kernel_inter_rank_gqa_fwd_batch_decode_combine_kv",
    "AllGatherMethod", "ag_gemm", "cp_engine_producer_all_gather_inter_node", "cp_engine_producer_all_gather_intra_node",
    "create_ag_gemm_context", "create_fast_allgather_context", "create_gemm_ar_context", "create_gemm_rs_context",
    "create_ll_gemm_ar_context", "fast_allgather", "gemm_allreduce_op", "gemm_non_persistent", "gemm_persistent", "gemm_rs",
    "get_auto_all_gather_method", "low_latency_gemm_allreduce_op", "_forward_pull_kernel", "_forward_push_2d_kernel",
    "_forward_push_2d_ll_kernel", "_forward_push_2d_ll_multimem_kernel", "_forward_push_3d_kernel",
    "_forward_push_numa_2d_kernel", "_forward_push_numa_2d_ll_kernel", "_forward_push_numa_2d_ll_multinode_kernel",
]
