"""Task record of the native schedule (reference core/task_base.py: TaskBase with layer_id / task_id / tile_id /
dependency / io tensors, serialised to INT_PER_TASK ints by core/scheduler.py:40-62)."""
from __future__ import annotations

import dataclasses
from typing import List

from .registry import op_name


@dataclasses.dataclass
class Task:
    op: int        # index into the op table
    type: int      # op code (registry.BY_CODE)
    tile: int      # tile id inside the op
    dep0: int      # scoreboard slot (= op index) to wait on, -1 = none
    dep0_cnt: int  # tiles of that op that must have completed
    dep1: int
    dep1_cnt: int
    out_slot: int  # scoreboard slot this task increments

    @property
    def type_name(self) -> str:
        return op_name(self.type)

    @staticmethod
    def from_ints(v: List[int]) -> "Task":
        return Task(*[int(x) for x in v[:8]])
