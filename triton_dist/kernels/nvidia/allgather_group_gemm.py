"""AllGather + grouped GEMM for tensor-parallel MoE (reference ``kernels/nvidia/allgather_group_gemm.py``:
context :199-318, ``ag_group_gemm`` :343-531, kernels K42-K44).

Every rank holds M/W tokens and a [E, N/W, K] shard of every expert's first FFN matrix. Steps on B200:
(1) low-latency push all-gather of the token shards into a symmetric buffer (one kernel, NVLink stores / NVLS),
(2) ``moe_align_sort`` of the gathered top-k ids (one single-CTA kernel, no host sync), (3) row gather into
expert-sorted 128-row blocks, (4) tcgen05 grouped GEMM with a per-block expert table. The output stays in
expert-sorted order for ``moe_reduce_rs``."""
from __future__ import annotations

import dataclasses
from typing import Optional

import torch
import torch.distributed as dist

from ...runtime import symm as _symm
from . import low_latency_allgather as _ag
from .moe_utils import MoESortResult, grouped_gemm, moe_align_sort, moe_gather_rows


@dataclasses.dataclass
class MoEAllGatherGroupGEMMTensorParallelContext:
    rank: int
    num_ranks: int
    num_local_ranks: int
    max_ntokens: int
    hidden: int
    num_experts: int
    topk: int
    dtype: torch.dtype
    symm_ag: Optional[torch.Tensor] = None
    ag_ctx: Optional[_ag.FastAllGatherContext] = None
    sort: Optional[MoESortResult] = None
    full_topk_ids: Optional[torch.Tensor] = None

    def __post_init__(self):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        self.symm_ag = heap.alloc((self.num_ranks, self.max_ntokens // self.num_ranks, self.hidden), self.dtype, zero=False)
        # per-rank id slot padded to a multiple of 16 bytes (the all-gather moves 16-byte vectors)
        self.ids_per_rank = (self.max_ntokens // self.num_ranks * self.topk + 3) // 4 * 4
        self.symm_ids = heap.alloc((self.num_ranks, self.ids_per_rank), torch.int32, zero=True)
        self.ag_ctx = _ag.create_fast_allgather_context(self.rank, 0, self.num_ranks, 1,
                                                        max_buffer_size=max(4 * self.symm_ag.numel() * self.symm_ag.element_size(), 1 << 20))

    def finalize(self):
        heap = _symm.get_heap()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.ag_ctx.finalize()
        heap.free(self.symm_ag)
        heap.free(self.symm_ids)


def create_ag_group_gemm_context(max_ntokens: int, hidden: int, num_experts: int, topk: int, rank: int, num_ranks: int,
                                 num_local_ranks: Optional[int] = None, dtype=torch.bfloat16, **_ignored):
    return MoEAllGatherGroupGEMMTensorParallelContext(rank=rank, num_ranks=num_ranks, num_local_ranks=num_local_ranks or num_ranks,
                                                      max_ntokens=max_ntokens, hidden=hidden, num_experts=num_experts, topk=topk,
                                                      dtype=dtype)


def ag_group_gemm(x: torch.Tensor, w: torch.Tensor, topk_ids: torch.Tensor, ctx: MoEAllGatherGroupGEMMTensorParallelContext):
    """x [M/W, K] local tokens, w [E, N/W, K] local expert shards, topk_ids [M/W, topk] local routing.
    Returns (y_sorted [max_rows, N/W], sort): expert-sorted activations of ALL M tokens and the sort layout."""
    W, r = ctx.num_ranks, ctx.rank
    m_loc = x.shape[0]
    assert m_loc * W <= ctx.max_ntokens and m_loc == ctx.max_ntokens // W, "token count must match the context (max_ntokens / W per rank)"
    if W > 1 and x.is_cuda:
        # symm_ag / symm_ids are single-buffered: no rank may push call N+1 into them while a slower rank still sorts or
        # gathers from call N (the reference opens the op the same way: barrier, copy, barrier — :44-81)
        _symm.get_heap().barrier_all()
    ctx.symm_ag[r].copy_(x)
    ctx.symm_ids[r, :m_loc * ctx.topk].copy_(topk_ids.to(torch.int32).reshape(-1))
    if W > 1:
        mode = "push_multimem" if x.is_cuda else "pull"
        _ag.fast_allgather(ctx.symm_ag, ctx.ag_ctx, mode=mode)
        _ag.fast_allgather(ctx.symm_ids, ctx.ag_ctx, mode=mode)
    full_x = ctx.symm_ag.view(W * m_loc, -1)
    full_ids = ctx.symm_ids[:, :m_loc * ctx.topk].reshape(W * m_loc, ctx.topk)
    sort = moe_align_sort(full_ids, ctx.num_experts)
    a_sorted = moe_gather_rows(full_x, sort)
    y = grouped_gemm(a_sorted, w, sort)
    ctx.sort, ctx.full_topk_ids = sort, full_ids
    return y, sort
