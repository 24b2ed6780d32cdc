"""Tile order for AllGather + MoE grouped GEMM (reference ``threadblock_swizzle_ag_moe.py`` / ``_triton.py:95-216`` and the
``.cu`` / ``.cc`` prototypes next to it).

With the tokens all-gathered shard by shard (rank r's own shard first, then r+1, ...), a 128-row block of the
expert-sorted layout can be computed as soon as the shards of all its rows have arrived. ``threadblock_swizzle_ag_moe``
returns, per block, the range of arrival stages it needs and a stable order of the blocks by their last stage, computed
by one native kernel (``moe_tile_swizzle_kernel``, csrc/kernels/moe.cu) without a host round trip.

Status on this build: the schedule is produced and tested; the grouped GEMM of ``ag_group_gemm`` still starts after the
NVLS all-gather has completed (the gather is a few microseconds per MB over NVSwitch), so the order is informational
there. It is what a stage-gated grouped GEMM would consume.
"""
from __future__ import annotations

import dataclasses

import torch

from ... import _lib
from .moe_utils import BLOCK_M, MoESortResult


@dataclasses.dataclass
class AgMoeTileSchedule:
    block_order: torch.Tensor  # [blocks] int32: block indices, ordered by the stage of their last-arriving shard
    segment_start: torch.Tensor  # [blocks] int32: first arrival stage a block needs
    segment_end: torch.Tensor  # [blocks] int32: last arrival stage a block needs
    stage_count: torch.Tensor  # [world] int32: number of blocks that become ready at each stage


def torch_threadblock_swizzle_ag_moe(sort: MoESortResult, tokens_per_rank: int, rank: int, world: int,
                                     block_m: int = BLOCK_M) -> AgMoeTileSchedule:
    ids = sort.sorted_pair_ids.long().cpu()
    nblk = int(sort.meta[1])
    seg_s, seg_e = torch.zeros(nblk, dtype=torch.int32), torch.zeros(nblk, dtype=torch.int32)
    for b in range(nblk):
        row = ids[b * block_m:(b + 1) * block_m]
        row = row[row < sort.num_pairs]
        if row.numel():
            st = ((row // sort.topk) // tokens_per_rank - rank) % world
            seg_s[b], seg_e[b] = int(st.min()), int(st.max())
    order = torch.argsort(seg_e, stable=True).to(torch.int32)
    counts = torch.bincount(seg_e.long(), minlength=world).to(torch.int32)
    dev = sort.sorted_pair_ids.device
    return AgMoeTileSchedule(order.to(dev), seg_s.to(dev), seg_e.to(dev), counts.to(dev))


def threadblock_swizzle_ag_moe(sort: MoESortResult, tokens_per_rank: int, rank: int, world: int,
                               block_m: int = BLOCK_M) -> AgMoeTileSchedule:
    """Device-side schedule (valid entries: the first ``sort.meta[1]`` blocks)."""
    if not sort.sorted_pair_ids.is_cuda:
        return torch_threadblock_swizzle_ag_moe(sort, tokens_per_rank, rank, world, block_m)
    dev = sort.sorted_pair_ids.device
    max_blocks = sort.max_rows // block_m
    res = AgMoeTileSchedule(torch.zeros(max_blocks, dtype=torch.int32, device=dev), torch.zeros(max_blocks, dtype=torch.int32, device=dev),
                            torch.zeros(max_blocks, dtype=torch.int32, device=dev), torch.zeros(_lib.MAX_RANKS, dtype=torch.int32, device=dev))
    _lib.check(_lib.lib().tdist_moe_tile_swizzle(sort.sorted_pair_ids.data_ptr(), sort.meta.data_ptr(), sort.num_pairs, sort.topk,
                                                 tokens_per_rank, rank, world, block_m, res.block_order.data_ptr(),
                                                 res.segment_start.data_ptr(), res.segment_end.data_ptr(), res.stage_count.data_ptr(),
                                                 max_blocks, _lib.stream_ptr()), "moe_tile_swizzle")
    res.stage_count = res.stage_count[:world]
    return res
