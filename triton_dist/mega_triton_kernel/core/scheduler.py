"""Static scheduler of the megakernel (reference core/scheduler.py:40-95: tasks are dealt round-robin, in op order, into
one work queue per SM). The dealing is done natively (csrc/mega/builder.cc); this module reads the result back and
restates the policy in Python for tests and documentation."""
from __future__ import annotations

import ctypes
from typing import List

from ... import _lib
from .task_base import Task


def round_robin_scheduler(tiles_per_op: List[int], num_ctas: int) -> List[List[tuple]]:
    """Pure-Python restatement: returns per CTA the list of (op, tile) in execution order."""
    queues: List[List[tuple]] = [[] for _ in range(num_ctas)]
    cursor = 0
    for op, n in enumerate(tiles_per_op):
        for t in range(n):
            queues[cursor % num_ctas].append((op, t))
            cursor += 1
    return queues


def get_queue(builder, cta: int) -> List[Task]:
    """The compiled queue of one CTA of a native ModelBuilder."""
    info = builder.info()
    cap = info["max_queue"]
    buf = (ctypes.c_int * (8 * cap))()
    n = _lib.lib().tdist_mega_get_queue(builder._handle, cta, buf, cap)
    return [Task.from_ints(buf[8 * i:8 * i + 8]) for i in range(max(n, 0))]


def describe(builder, max_ctas: int = 2) -> str:
    """Human-readable dump of the first queues (what the reference's generated kernel source shows)."""
    lines = [f"megakernel graph: {builder.info()}"]
    if builder.native:
        for c in range(min(max_ctas, builder.info()["ctas"])):
            q = get_queue(builder, c)
            lines.append(f"CTA {c}: " + " -> ".join(f"{t.type_name}#{t.op}[{t.tile}]" for t in q[:24]) + (" ..." if len(q) > 24 else ""))
    return "\n".join(lines)
