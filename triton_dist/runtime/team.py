"""Teams: ordered subsets of the ranks of the symmetric heap (the reference reaches NVSHMEM's teams through nvshmem4py:
``team_split_strided`` / ``team_split_2d`` / ``team_my_pe`` / ``team_n_pes`` / ``team_translate_pe``, exercised by
test/nvidia/test_team_split.py to build EP x PP groups).

The heap of this framework is always mapped over the whole world (one NVSwitch node), so a team is pure rank arithmetic
plus a barrier over its members: no second heap, no extra mappings. Kernels take the world ``SymmCtx``; code that works
on a sub-group translates team ranks to world ranks with ``Team.world_rank`` before calling ``symm_at`` / ``peer_view``.
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional, Tuple

import torch


@dataclasses.dataclass(frozen=True)
class Team:
    start: int   # world rank of team rank 0
    stride: int  # world-rank distance between consecutive team ranks
    size: int
    world_rank_self: int  # world rank of the calling process

    # ------------------------------------------------------------------ queries (nvshmem_team_my_pe / n_pes)
    def ranks(self) -> List[int]:
        return [self.start + i * self.stride for i in range(self.size)]

    def n_pes(self) -> int:
        return self.size

    def my_pe(self) -> int:
        """Rank of the caller inside the team, -1 when it is not a member."""
        d = self.world_rank_self - self.start
        if d < 0 or d % self.stride != 0 or d // self.stride >= self.size:
            return -1
        return d // self.stride

    def world_rank(self, team_rank: int) -> int:
        assert 0 <= team_rank < self.size
        return self.start + team_rank * self.stride

    def translate_pe(self, team_rank: int, dest: "Team") -> int:
        """Rank in ``dest`` of the process that is ``team_rank`` here, -1 if it is not in ``dest`` (team_translate_pe)."""
        w = self.world_rank(team_rank)
        d = w - dest.start
        if d < 0 or d % dest.stride != 0 or d // dest.stride >= dest.size:
            return -1
        return d // dest.stride

    # ------------------------------------------------------------------ splits
    def split_strided(self, start: int, stride: int, size: int) -> "Team":
        """Sub-team of team ranks start, start + stride, ... (nvshmem_team_split_strided)."""
        assert size >= 1 and stride >= 1 and start + (size - 1) * stride < self.size
        return Team(self.world_rank(start), self.stride * stride, size, self.world_rank_self)

    def split_2d(self, xrange: int) -> Tuple["Team", "Team"]:
        """(x team, y team) of the caller for a row-major [size / xrange, xrange] grid (nvshmem_team_split_2d): the x team
        holds the ``xrange`` consecutive ranks of my row, the y team the ranks of my column."""
        assert self.size % xrange == 0
        me = self.my_pe()
        assert me >= 0, "the caller is not a member of the parent team"
        row, col = divmod(me, xrange)
        x = Team(self.world_rank(row * xrange), self.stride, xrange, self.world_rank_self)
        y = Team(self.world_rank(col), self.stride * xrange, self.size // xrange, self.world_rank_self)
        return x, y

    # ------------------------------------------------------------------ synchronisation
    def barrier(self, flags: torch.Tensor, epoch: int, stream: Optional["torch.cuda.Stream"] = None) -> None:
        """Barrier over the members on ``stream``: ``flags`` is a symmetric int32 tensor with at least ``size`` entries
        that belongs to this team, ``epoch`` a value that grows with every call (monotone flags, nothing is reset)."""
        from ..kernels.common_ops import CMP_GE, set_signal, wait_eq
        from .symm import get_heap

        me = self.my_pe()
        if me < 0:
            return
        heap = get_heap()
        if not flags.is_cuda:  # host heap: the members meet through torch.distributed
            import torch.distributed as dist

            if dist.is_initialized() and self.size > 1:
                dist.barrier()  # (world barrier: correct, just stronger than needed on the CPU plumbing path)
            return
        for r in range(self.size):
            set_signal(heap.peer_view(flags, self.world_rank(r))[me:me + 1], epoch, stream)
        for r in range(self.size):
            wait_eq(flags[r:r + 1], epoch, stream, cmp=CMP_GE)


def team_world(world_size: Optional[int] = None, rank: Optional[int] = None) -> Team:
    import torch.distributed as dist

    if world_size is None:
        world_size = dist.get_world_size() if dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    return Team(0, 1, world_size, rank)
