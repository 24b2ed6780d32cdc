"""Tile order of the GEMM+RS kernel (reference ``gemm_rs_threadblock_swizzle.py:68-153``).

Like AG+GEMM but the destination segments are visited starting at rank + 1 and ending with the caller's own, so that
the partial tiles other ranks are waiting for are produced first (gemm_reduce_scatter.py:188,198)."""
from __future__ import annotations

from typing import List, Tuple


def gemm_rs_tile_order(M: int, N: int, world: int, rank: int, block_m: int = 128, block_n: int = 256, group_m: int = 8
                       ) -> List[Tuple[int, int, int]]:
    rows = M // world
    bps = (rows + block_m - 1) // block_m
    tiles_n = (N + block_n - 1) // block_n
    out = []
    for step in range(world):
        shard = (rank + step + 1) % world
        for t in range(bps * tiles_n):
            group = t // (group_m * tiles_n)
            first = group * group_m
            size = min(group_m, bps - first)
            local = t - group * group_m * tiles_n
            out.append((shard, first + local % size, local // size))
    return out


def threadblock_swizzle_gemm_reduce_scatter(tile_id: int, M: int, N: int, world: int, rank: int, **kw) -> Tuple[int, int, int]:
    return gemm_rs_tile_order(M, N, world, rank, **kw)[tile_id]
