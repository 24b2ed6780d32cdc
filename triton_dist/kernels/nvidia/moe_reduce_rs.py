"""Grouped GEMM + top-k reduce + reduce-scatter for tensor-parallel MoE (reference
``kernels/nvidia/moe_reduce_rs.py``: context :60-160, grouped GEMM K46 :166-244, fused reduce/scatter K47 :292-376,
``run_moe_reduce_rs`` :528-660).

Second FFN matrix of every expert is sharded on its input dim ([E, H, I/W]). Steps on B200: tcgen05 grouped GEMM on
the expert-sorted activations, then ONE kernel that applies the top-k weights, sums the top-k rows of every token and
pushes each token segment to its owner rank over NVLink, followed (same kernel) by the owner's fp32 sum over ranks."""
from __future__ import annotations

import ctypes
import dataclasses
from typing import Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ..common_ops import CollectiveState
from .moe_utils import MoESortResult, grouped_gemm, reduce_topk
from .reduce_scatter import torch_reduce_scatter


@dataclasses.dataclass
class MoEReduceRSContext:
    rank: int
    world_size: int
    local_world_size: int
    max_token_num: int
    hidden_dim: int
    num_experts: int
    topk: int
    input_dtype: torch.dtype
    output_dtype: torch.dtype
    n_chunks_max: int = 4
    scatter_buf: Optional[torch.Tensor] = None
    state: Optional[CollectiveState] = None

    def __post_init__(self):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        if self.native:
            self.scatter_buf = heap.alloc((2, self.world_size, self.max_token_num // self.world_size, self.hidden_dim),
                                          self.output_dtype, zero=False)
            self.state = CollectiveState()
            heap.barrier_all()

    def finalize(self):
        if self.native and self.scatter_buf is not None:
            torch.cuda.synchronize()
            _symm.get_heap().free(self.scatter_buf)
            self.state.free()
            self.scatter_buf = None


def create_moe_rs_context(rank, world_size, local_world_size, max_token_num, hidden_dim, num_experts, topk, input_dtype,
                          output_dtype=None, device=None, moe_block_size=128, n_chunks_max=4, **_ignored) -> MoEReduceRSContext:
    return MoEReduceRSContext(rank=rank, world_size=world_size, local_world_size=local_world_size, max_token_num=max_token_num,
                              hidden_dim=hidden_dim, num_experts=num_experts, topk=topk, input_dtype=input_dtype,
                              output_dtype=output_dtype or input_dtype, n_chunks_max=n_chunks_max)


def run_moe_reduce_rs(h_sorted: torch.Tensor, w: torch.Tensor, sort: MoESortResult, topk_weights: torch.Tensor,
                      ctx: MoEReduceRSContext, n_chunks: int = 4) -> torch.Tensor:
    """h_sorted [max_rows, I/W] (expert-sorted), w [E, H, I/W], topk_weights [M, topk] for ALL tokens.
    Returns this rank's token segment [M/W, H]."""
    y = grouped_gemm(h_sorted, w, sort)  # [max_rows, H] partial over the FFN shard
    tokens = sort.num_pairs // sort.topk
    W = ctx.world_size
    wts = topk_weights.float().contiguous()
    if not ctx.native:
        full = reduce_topk(y, sort, wts)
        return torch_reduce_scatter(full) if W > 1 else full
    out = torch.empty((tokens // W, y.shape[1]), dtype=ctx.output_dtype, device=y.device)
    heap = _symm.get_heap()
    st = ctx.state
    _lib.check(_lib.lib().tdist_moe_reduce_rs(ctypes.byref(heap.ctx), _lib.dtype_code(y.dtype), y.data_ptr(),
                                              ctx.scatter_buf.data_ptr(), ctx.scatter_buf.numel() * ctx.scatter_buf.element_size(),
                                              out.data_ptr(), sort.pair_to_row.data_ptr(),
                                              wts.data_ptr(), tokens, sort.topk, y.shape[1] * y.element_size(), st.flags_ptr,
                                              st.phase_ptr, st.done_ptr, 64, _lib.stream_ptr()), "moe_reduce_rs")
    return out
