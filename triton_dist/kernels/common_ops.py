"""Host-side stream signalling and barrier helpers.

Reference: ``kernels/common_ops.py:25-39`` (``wait_eq`` / ``set_signal``) and
``kernels/nvidia/common_ops.py:212-397`` (``BarrierAllContext``,
``barrier_all_on_stream``, cuStreamWriteValue/WaitValue wrappers).

On B200 the signal helpers default to one-thread *kernels* so that they can be
captured in CUDA graphs; ``use_memops=True`` selects cuStreamWriteValue32 /
cuStreamWaitValue32 (what copy-engine producers use).
"""
from __future__ import annotations

import ctypes
import dataclasses
from typing import Optional

import torch
import torch.distributed as dist

from .. import _lib
from ..runtime import symm as _symm

CMP_EQ, CMP_NE, CMP_GT, CMP_LE, CMP_LT, CMP_GE = range(6)


def set_signal(signal: torch.Tensor, value: int, stream: Optional["torch.cuda.Stream"] = None, add: bool = False,
               use_memops: bool = False) -> None:
    """Write (or add) ``value`` into a 32-bit flag in stream order."""
    if not signal.is_cuda:
        if add:
            signal.add_(value)
        else:
            signal.fill_(value)
        return
    lib = _lib.lib()
    if use_memops and not add:
        _lib.check(lib.tdist_stream_write_value32(_lib.stream_ptr(stream), signal.data_ptr(), value), "stream write value")
    else:
        _lib.check(lib.tdist_signal_set(signal.data_ptr(), value, int(add), _lib.stream_ptr(stream)), "signal_set")


def wait_eq(signal: torch.Tensor, value: int, stream: Optional["torch.cuda.Stream"] = None, use_memops: bool = False,
            cmp: int = CMP_EQ) -> None:
    """Block the stream until the flag compares true against ``value``."""
    if not signal.is_cuda:
        return
    lib = _lib.lib()
    if use_memops and cmp in (CMP_EQ, CMP_GE):
        _lib.check(lib.tdist_stream_wait_value32(_lib.stream_ptr(stream), signal.data_ptr(), value, int(cmp == CMP_GE)),
                   "stream wait value")
    else:
        _lib.check(lib.tdist_signal_wait(signal.data_ptr(), value, cmp, _lib.stream_ptr(stream)), "signal_wait")


class CollectiveState:
    """Flag rows + monotone phase counter shared by the launches of one context.

    ``flags`` is symmetric (peers write into it), ``state`` = (phase, done) is
    only touched locally but lives in the heap as well so that contexts never
    call the caching allocator after creation.
    """

    def __init__(self, n_flags: int = _lib.MAX_CTAS * _lib.MAX_RANKS):
        heap = _symm.get_heap()
        self.heap = heap
        self.flags = heap.alloc((n_flags,), torch.int32)
        self.state = heap.alloc((2,), torch.int32)

    @property
    def flags_ptr(self) -> int:
        return self.flags.data_ptr()

    @property
    def phase_ptr(self) -> int:
        return self.state.data_ptr()

    @property
    def done_ptr(self) -> int:
        return self.state.data_ptr() + 4

    def free(self) -> None:
        self.heap.free(self.flags)
        self.heap.free(self.state)


@dataclasses.dataclass
class BarrierAllContext:
    """Context of the stream barrier (reference common_ops.py:212-229)."""

    is_intra_node: bool = True
    state: Optional[CollectiveState] = None

    def __post_init__(self):
        heap = _symm.get_heap()
        self.rank, self.world = heap.rank, heap.world
        if isinstance(heap, _symm.SymmetricHeap):
            self.state = CollectiveState(_lib.MAX_RANKS)
            heap.barrier_all()
            torch.cuda.synchronize()


def barrier_all_on_stream(ctx: Optional[BarrierAllContext] = None, stream: Optional["torch.cuda.Stream"] = None) -> None:
    """Cross-rank barrier enqueued on ``stream`` (graph-capturable)."""
    heap = _symm.get_heap()
    if not isinstance(heap, _symm.SymmetricHeap):
        heap.barrier_all()
        return
    if ctx is None or ctx.state is None:
        heap.barrier_all(stream)
        return
    st = ctx.state
    _lib.check(_lib.lib().tdist_barrier_all(ctypes.byref(heap.ctx), st.flags_ptr, st.phase_ptr, st.done_ptr,
                                            _lib.stream_ptr(stream)), "barrier_all")
