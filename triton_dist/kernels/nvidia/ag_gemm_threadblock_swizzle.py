"""Tile order of the AG+GEMM kernel (reference ``ag_gemm_threadblock_swizzle.py:51-137``: there a device function that
remaps tile-M by node / rank with warp shuffles when nnodes > 1).

Single NVSwitch node here: the order is the one `tile_coord` of csrc/kernels/gemm_sm100.cuh implements — shards in
arrival order starting with the caller's own, and inside a shard the L2-friendly group_m swizzle. This module restates it
on the host for inspection, tests and docs."""
from __future__ import annotations

from typing import List, Tuple


def ag_gemm_tile_order(M: int, N: int, world: int, rank: int, block_m: int = 128, block_n: int = 256, group_m: int = 8
                       ) -> List[Tuple[int, int, int]]:
    """Returns [(shard, m_block_in_shard, n_block)] in the order the persistent kernel walks the tiles."""
    rows = M // world
    bps = (rows + block_m - 1) // block_m
    tiles_n = (N + block_n - 1) // block_n
    out = []
    for step in range(world):
        shard = (rank + step) % world
        for t in range(bps * tiles_n):
            group = t // (group_m * tiles_n)
            first = group * group_m
            size = min(group_m, bps - first)
            local = t - group * group_m * tiles_n
            out.append((shard, first + local % size, local // size))
    return out


def threadblock_swizzle_allgather_gemm(tile_id: int, M: int, N: int, world: int, rank: int, **kw) -> Tuple[int, int, int]:
    return ag_gemm_tile_order(M, N, world, rank, **kw)[tile_id]
