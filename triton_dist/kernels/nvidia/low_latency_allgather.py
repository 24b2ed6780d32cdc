"""Low-latency all-gather of a symmetric buffer. API of the reference's
``kernels/nvidia/low_latency_allgather.py`` (kernels :47-777, context :781-816,
``fast_allgather`` + per-mode entry points :819-992).

``symm_buffer`` is a symmetric tensor of ``num_ranks`` equal slots; every rank
has filled slot ``rank`` and all slots are valid on return.

NVSwitch makes the node flat, so the reference's node x NUMA x GPU hierarchies
(2-D / 3-D / NUMA pushes) all resolve to a one-hop exchange here:
  pull                 barrier, then 16-byte loads from each peer's slot
  push_2d / push_3d / push_numa_2d   16-byte stores into every peer + flag barrier
  push_2d_ll / push_numa_2d_ll[_multinode]   LL packets {data, flag} (one NVLink trip, no fence)
  push_2d_ll_multimem  LL packets sent once with NVLS multimem.st
Kernels: csrc/kernels/collectives.cu (allgather_kernel, allgather_ll_kernel).
"""
from __future__ import annotations

import ctypes
import dataclasses
from typing import List, Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ...utils import launch_metadata
from ..common_ops import CollectiveState

AG_PULL, AG_PUSH, AG_PUSH_MC, AG_LL, AG_LL_MC = range(5)


@dataclasses.dataclass
class FastAllGatherContext:
    rank: int
    node: int
    num_ranks: int
    num_nodes: int
    max_buffer_size: int = 2 * 32 * 1024 * 1024
    signal_target: int = 1
    ll_buffer: Optional[torch.Tensor] = None
    state: Optional[CollectiveState] = None
    max_ctas: int = 16

    def __post_init__(self):
        if self.num_nodes != 1:
            raise NotImplementedError("single NVSwitch node only")
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        if self.native:
            # [2 parity][payload * 2] LL packets
            self.ll_buffer = heap.alloc((2 * self.max_buffer_size,), torch.int8, zero=True)
            self.state = CollectiveState()
            self.multimem = bool(heap.multicast)

    def finalize(self):
        if self.native and self.ll_buffer is not None:
            torch.cuda.synchronize()
            _symm.get_heap().free(self.ll_buffer)
            self.state.free()
            self.ll_buffer = None


def create_fast_allgather_context(rank, node, num_ranks, num_nodes, max_buffer_size: int = 2 * 32 * 1024 * 1024):
    ctx = FastAllGatherContext(rank=rank, node=node, num_ranks=num_ranks, num_nodes=num_nodes, max_buffer_size=max_buffer_size)
    _symm.get_heap().barrier_all()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    return ctx


def _torch_inplace_allgather(symm_buffer: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    if world > 1:
        flat = symm_buffer.view(world, -1)
        dist.all_gather(list(flat.unbind(0)), flat[rank].clone())
    return symm_buffer


def _launch(ctx: FastAllGatherContext, symm_buffer: torch.Tensor, mode: int) -> torch.Tensor:
    if not ctx.native:
        return _torch_inplace_allgather(symm_buffer, ctx.rank, ctx.num_ranks)
    heap = _symm.get_heap()
    assert heap.is_symmetric(symm_buffer), "fast_allgather needs a symmetric buffer"
    nbytes = symm_buffer.numel() * symm_buffer.element_size()
    assert nbytes % ctx.num_ranks == 0
    shard = nbytes // ctx.num_ranks
    base = symm_buffer.data_ptr()
    mine = base + ctx.rank * shard
    st = ctx.state
    if mode in (AG_PUSH_MC, AG_LL_MC) and not ctx.multimem:
        mode = AG_PUSH if mode == AG_PUSH_MC else AG_LL
    ctx.signal_target += 1
    if mode in (AG_LL, AG_LL_MC):
        assert nbytes * 2 <= ctx.max_buffer_size, f"LL all-gather payload {nbytes} B exceeds max_buffer_size/2"
        # packets: two parity halves of ll_buffer, each [W][shard/8] x 16 B; the halves are half the buffer apart
        # whatever the size of this call (a fast rank's next call must not land on packets a slow rank still reads)
        rc = _lib.lib().tdist_allgather(ctypes.byref(heap.ctx), mode, mine, ctx.ll_buffer.data_ptr(), base, shard,
                                        ctx.ll_buffer.numel() * ctx.ll_buffer.element_size(), st.flags_ptr, st.phase_ptr, st.done_ptr, ctx.max_ctas, _lib.stream_ptr())
    else:
        rc = _lib.lib().tdist_allgather(ctypes.byref(heap.ctx), mode, mine, base, None, shard, 0, st.flags_ptr, st.phase_ptr,
                                        st.done_ptr, ctx.max_ctas, _lib.stream_ptr())
    _lib.check(rc, "fast_allgather")
    return symm_buffer


def fast_allgather_pull(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_PULL)


def fast_allgather_push_2d(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_PUSH)


def fast_allgather_push_3d(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_PUSH)


def fast_allgather_push_2d_ll(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_LL)


def fast_allgather_push_2d_ll_multimem(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_LL_MC)


def fast_allgather_push_numa_2d(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_PUSH)


def fast_allgather_push_numa_2d_ll(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_LL)


def fast_allgather_push_numa_2d_ll_multinode(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_LL)


def fast_allgather_push_multimem(ctx, symm_buffer):
    return _launch(ctx, symm_buffer, AG_PUSH_MC)


FAST_ALLGATHER_FUNC_DISPATCH = {
    "pull": fast_allgather_pull,
    "push2d": fast_allgather_push_2d,
    "push3d": fast_allgather_push_3d,
    "push2d_ll": fast_allgather_push_2d_ll,
    "push2d_ll_multimem": fast_allgather_push_2d_ll_multimem,
    "push_numa_2d": fast_allgather_push_numa_2d,
    "push_numa_2d_ll": fast_allgather_push_numa_2d_ll,
    "push_numa_2d_ll_multinode": fast_allgather_push_numa_2d_ll_multinode,
    "push_multimem": fast_allgather_push_multimem,
}

# The reference exports its Triton kernel objects under these names; here they are the launchers.
_forward_pull_kernel = fast_allgather_pull
_forward_push_2d_kernel = fast_allgather_push_2d
_forward_push_3d_kernel = fast_allgather_push_3d
_forward_push_2d_ll_kernel = fast_allgather_push_2d_ll
_forward_push_2d_ll_multimem_kernel = fast_allgather_push_2d_ll_multimem
_forward_push_numa_2d_kernel = fast_allgather_push_numa_2d
_forward_push_numa_2d_ll_kernel = fast_allgather_push_numa_2d_ll
_forward_push_numa_2d_ll_multinode_kernel = fast_allgather_push_numa_2d_ll_multinode


@launch_metadata("fast_allgather", lambda buf, *_, mode="pull", **__: {"mode": mode, "bytes": buf.numel() * buf.element_size()})
def fast_allgather(symm_buffer: torch.Tensor, ctx: Optional[FastAllGatherContext] = None, rank=None, node=None,
                   num_ranks=None, num_nodes=None, mode: str = "pull") -> torch.Tensor:
    assert mode in FAST_ALLGATHER_FUNC_DISPATCH, f"unknown mode {mode}"
    if ctx is None:
        assert None not in (rank, node, num_ranks, num_nodes)
        ctx = create_fast_allgather_context(rank, node, num_ranks, num_nodes)
    return FAST_ALLGATHER_FUNC_DISPATCH[mode](ctx, symm_buffer)
