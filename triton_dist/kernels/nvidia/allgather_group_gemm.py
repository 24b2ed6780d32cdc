"""AllGather + grouped GEMM for tensor-parallel MoE (reference ``kernels/nvidia/allgather_group_gemm.py``:
context :199-318, ``ag_group_gemm`` :343-531, kernels K42-K44).

Every rank holds M/W tokens and a [E, N/W, K] shard of every expert's first FFN matrix. Two schedules:

* default: (1) low-latency push all-gather of the token shards into a symmetric buffer (one kernel, NVLink stores /
  NVLS), (2) ``moe_align_sort`` of the gathered top-k ids (one single-CTA kernel, no host sync), (3) row gather into
  expert-sorted 128-row blocks, (4) tcgen05 grouped GEMM with a per-block expert table.
* ``overlap=True`` (context argument, or ``TDIST_AG_MOE_OVERLAP=1``): the reference's schedule (:414-490 copy-engine
  producer with per-source signals, :534-608 consumer tiles that wait for the source ranks they need). Only the routing
  ids are gathered up front (a few KB); the token shards are PULLED by the copy engines on a side stream, own rank + 1
  first, each followed by a stream-ordered signal. ``threadblock_swizzle_ag_moe`` tells, for every 128-row block of the
  sorted layout, at which arrival stage its last token is there; the compute stream then runs, for stage 0 .. W-1:
  wait for that stage's signal (a one-thread wait kernel in stream order: it waits for a DMA, never for another kernel, and the
  host is not involved), gather the rows of the blocks that
  became complete, grouped GEMM restricted to those blocks (the others carry expert -1 and are skipped on the device).
  The transfer of shard s+1 overlaps gather + GEMM of stage s and uses no SM. Built from the same kernels as the default
  schedule plus a stage-filtered gather; numerics are identical by construction. It is opt-in because it has not been
  timed on hardware yet.

The output stays in expert-sorted order for ``moe_reduce_rs``."""
from __future__ import annotations

import dataclasses
from typing import Optional

import torch
import torch.distributed as dist

import os

from ...runtime import symm as _symm
from ...utils import gemm_metadata, launch_metadata
from ..common_ops import CMP_GE, wait_eq
from . import low_latency_allgather as _ag
from .allgather import cp_engine_producer_all_gather_intra_node
from .moe_utils import MoESortResult, grouped_gemm, moe_align_sort, moe_gather_rows, moe_gather_rows_stage
from .threadblock_swizzle_ag_moe import threadblock_swizzle_ag_moe


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
    overlap: bool = False

    def __post_init__(self):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        self.symm_ag = heap.alloc((self.num_ranks, self.max_ntokens // self.num_ranks, self.hidden), self.dtype, zero=False)
        # per-rank id slot padded to a multiple of 16 bytes (the all-gather moves 16-byte vectors)
        self.ids_per_rank = (self.max_ntokens // self.num_ranks * self.topk + 3) // 4 * 4
        self.symm_ids = heap.alloc((self.num_ranks, self.ids_per_rank), torch.int32, zero=True)
        self.ag_ctx = _ag.create_fast_allgather_context(self.rank, 0, self.num_ranks, 1,
                                                        max_buffer_size=max(4 * self.symm_ag.numel() * self.symm_ag.element_size(), 1 << 20))
        self.overlap = self.overlap or os.environ.get("TDIST_AG_MOE_OVERLAP", "0") == "1"
        self.symm_sig, self.ag_stream, self.call = None, None, 0
        if self.overlap:
            # stage-gated schedule: per-source arrival signals (monotone call counter, never reset) + the copy-engine stream
            self.symm_sig = heap.alloc((self.num_ranks,), torch.int32, zero=True)
            if self.native:
                self.ag_stream = torch.cuda.Stream()
                self.ag_views = heap.peer_views(self.symm_ag.view(self.max_ntokens, self.hidden))
                self.sig_views = heap.peer_views(self.symm_sig)
                heap.barrier_all()  # every rank's signals are zero before the first producer writes one

    def finalize(self):
        heap = _symm.get_heap()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.ag_ctx.finalize()
        if self.ag_stream is not None:
            self.ag_stream.synchronize()
        heap.free(self.symm_ag)
        heap.free(self.symm_ids)
        if self.symm_sig is not None:
            heap.free(self.symm_sig)


def create_ag_group_gemm_context(max_ntokens: int, hidden: int, num_experts: int, topk: int, rank: int, num_ranks: int,
                                 num_local_ranks: Optional[int] = None, dtype=torch.bfloat16, overlap: bool = False, **_ignored):
    return MoEAllGatherGroupGEMMTensorParallelContext(rank=rank, num_ranks=num_ranks, num_local_ranks=num_local_ranks or num_ranks,
                                                      max_ntokens=max_ntokens, hidden=hidden, num_experts=num_experts, topk=topk,
                                                      dtype=dtype, overlap=overlap)


@launch_metadata("ag_group_gemm", lambda x, w, ids, ctx, *_, **__: gemm_metadata(x.shape[0] * ctx.num_ranks * ctx.topk, w.shape[1], w.shape[2], x.element_size(),
                                                                                 ctx.num_ranks, (ctx.num_ranks - 1) * x.numel() * x.element_size()))
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
    mode = "push_multimem" if x.is_cuda else "pull"
    if ctx.overlap and W > 1:
        return _ag_group_gemm_staged(x, w, ctx, m_loc, mode)
    if W > 1:
        _ag.fast_allgather(ctx.symm_ag, ctx.ag_ctx, mode=mode)
        _ag.fast_allgather(ctx.symm_ids, ctx.ag_ctx, mode=mode)
    full_x = ctx.symm_ag.view(W * m_loc, -1)
    full_ids = ctx.symm_ids[:, :m_loc * ctx.topk].reshape(W * m_loc, ctx.topk)
    sort = moe_align_sort(full_ids, ctx.num_experts)
    a_sorted = moe_gather_rows(full_x, sort)
    y = grouped_gemm(a_sorted, w, sort)
    ctx.sort, ctx.full_topk_ids = sort, full_ids
    return y, sort


def _ag_group_gemm_staged(x, w, ctx: MoEAllGatherGroupGEMMTensorParallelContext, m_loc: int, mode: str):
    """The stage-gated schedule (see the module docstring). Called with this rank's tokens and ids already in its own
    slots of the symmetric buffers."""
    W, r = ctx.num_ranks, ctx.rank
    # routing first: a few KB. The kernel ends with a cross-rank exchange, so once it has completed every rank has (in
    # stream order) filled its slot of symm_ag too: the pulls below read complete shards.
    _ag.fast_allgather(ctx.symm_ids, ctx.ag_ctx, mode=mode)
    full_x = ctx.symm_ag.view(W * m_loc, -1)
    full_ids = ctx.symm_ids[:, :m_loc * ctx.topk].reshape(W * m_loc, ctx.topk)
    ctx.call += 1
    if ctx.native:
        # producer: copy engines pull shard (r+1), (r+2), ... and write signal[src] = call after each one
        ctx.ag_stream.wait_stream(torch.cuda.current_stream())
        cp_engine_producer_all_gather_intra_node(r, W, x, ctx.ag_views, ctx.ag_stream, ctx.sig_views, signal_target=ctx.call)
    else:
        _ag.fast_allgather(ctx.symm_ag, ctx.ag_ctx, mode=mode)  # CPU plumbing: no streams, the stage loop still runs
    sort = moe_align_sort(full_ids, ctx.num_experts)
    sched = threadblock_swizzle_ag_moe(sort, m_loc, r, W)
    n_blocks = sort.block_expert.numel()
    stage_of_block = torch.zeros(n_blocks, dtype=torch.int32, device=x.device)  # unused blocks: expert -1, any stage
    n_sched = min(n_blocks, sched.segment_end.numel())
    stage_of_block[:n_sched] = sched.segment_end[:n_sched]
    a_sorted = torch.empty((sort.max_rows, full_x.shape[1]), dtype=x.dtype, device=x.device)
    y = torch.empty((sort.max_rows, w.shape[1]), dtype=x.dtype, device=x.device)
    skip = torch.full_like(sort.block_expert, -1)
    for stage in range(W):
        if stage > 0 and ctx.native:
            src = (r + stage) % W
            wait_eq(ctx.symm_sig[src:src + 1], ctx.call, cmp=CMP_GE)  # the compute STREAM (not the host) waits for shard `src`
        moe_gather_rows_stage(full_x, sort, stage_of_block, stage, a_sorted)
        grouped_gemm(a_sorted, w, sort, out=y, block_expert=torch.where(stage_of_block == stage, sort.block_expert, skip))
    if ctx.native:
        torch.cuda.current_stream().wait_stream(ctx.ag_stream)
    ctx.sort, ctx.full_topk_ids = sort, full_ids
    return y, sort
