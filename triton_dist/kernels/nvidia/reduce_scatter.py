"""Reduce-scatter over the symmetric heap. API of the reference's
``kernels/nvidia/reduce_scatter.py`` (context :44-179, ``reduce_scatter_2d_op``
:857-866, ``ring_reduce`` :815-825).

B200 implementation: one kernel (csrc/kernels/collectives.cu,
``reduce_scatter_kernel``). Every rank stores segment ``s`` of its input
directly into slot ``rank`` of rank ``s``'s scatter buffer over NVLink, a
per-CTA flag exchange tells the owner when a slice from all ranks has landed,
and the owner sums the W slots in fp32. The ring / 2-D / inter-node variants of
the reference exist for PCIe and multi-node boxes and are not needed behind
NVSwitch; their entry points are kept and resolve to the same kernel.
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
from ..common_ops import BarrierAllContext, CollectiveState


@dataclasses.dataclass
class ReduceScatter2DContext:
    max_M: int
    N: int
    rank: int
    world_size: int
    local_world_size: int
    dtype: torch.dtype
    overlap_with_gemm: bool = True
    num_reduction_sms: int = 32
    scatter_buf: Optional[torch.Tensor] = dataclasses.field(default=None, init=False)
    scatter_bufs: List[torch.Tensor] = dataclasses.field(default_factory=list, init=False)
    state: Optional[CollectiveState] = dataclasses.field(default=None, init=False)
    barrier: Optional[BarrierAllContext] = dataclasses.field(default=None, init=False)

    def __post_init__(self):
        assert self.world_size % self.local_world_size == 0
        assert self.max_M % self.world_size == 0, "max_M must be divisible by world_size"
        self.local_rank = self.rank % self.local_world_size
        self.node_id = self.rank // self.local_world_size
        self.nnodes = self.world_size // self.local_world_size
        if self.nnodes != 1:
            raise NotImplementedError("single NVSwitch node only")
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        if self.native:
            # [2 parity][W slots][max_M / W][N]
            self.scatter_buf = heap.alloc((2, self.world_size, self.max_M // self.world_size, self.N), self.dtype, zero=False)
            self.scatter_bufs = heap.peer_views(self.scatter_buf)
            # fused GEMM+RS, NVLS protocol: one flag per (source rank, 128-row block, n tile >= 32 wide); the stand-alone
            # reduce-scatter needs a W-wide row per CTA
            blocks = self.world_size * ((self.max_M // self.world_size + 127) // 128)
            self.state = CollectiveState(max(_lib.MAX_CTAS * _lib.MAX_RANKS, blocks * ((self.N + 31) // 32) + 64))
            self.barrier = BarrierAllContext(True)
            self.reduction_stream = torch.cuda.Stream(priority=-1)

    @property
    def num_rs_sms(self) -> int:
        return 0  # the scatter rides in the GEMM epilogue; no SMs are reserved

    def finalize(self):
        if self.native and self.scatter_buf is not None:
            torch.cuda.synchronize()
            heap = _symm.get_heap()
            heap.free(self.scatter_buf)
            self.state.free()
            self.scatter_buf = None


def create_reduce_scater_2d_ctx(max_M, N, rank, world_size, local_world_size, dtype, overlap_with_gemm=True,
                                num_reduction_sms=32) -> ReduceScatter2DContext:
    """(sic) spelling kept from the reference, reduce_scatter.py:148."""
    ctx = ReduceScatter2DContext(max_M=max_M, N=N, rank=rank, world_size=world_size,
                                 local_world_size=min(local_world_size, world_size), dtype=dtype,
                                 overlap_with_gemm=overlap_with_gemm, num_reduction_sms=num_reduction_sms)
    _symm.get_heap().barrier_all()
    return ctx


create_reduce_scatter_2d_ctx = create_reduce_scater_2d_ctx


def torch_reduce_scatter(x: torch.Tensor, group=None) -> torch.Tensor:
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return x.clone()
    out = torch.empty((x.shape[0] // world,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    if x.is_cuda:
        dist.reduce_scatter_tensor(out, x.contiguous(), group=group)
    else:  # gloo has no reduce_scatter: all_reduce + slice
        full = x.clone()
        dist.all_reduce(full, group=group)
        r = dist.get_rank(group)
        out.copy_(full[r * out.shape[0]:(r + 1) * out.shape[0]])
    return out


@launch_metadata("reduce_scatter", lambda x, ctx, *_, **__: {"M": x.shape[0], "N": x.shape[1], "W": ctx.world_size, "bytes": x.numel() * x.element_size()})
def reduce_scatter_2d_op(input: torch.Tensor, ctx: ReduceScatter2DContext, output: Optional[torch.Tensor] = None) -> torch.Tensor:
    """output[M/W, N] = sum over ranks of input[rank*M/W:(rank+1)*M/W] (input: [M, N] per rank)."""
    M, N = input.shape
    W = ctx.world_size
    assert M % W == 0 and M <= ctx.max_M and N == ctx.N
    if not ctx.native:
        res = torch_reduce_scatter(input)
        if output is not None:
            output.copy_(res)
            return output
        return res
    if output is None:
        output = torch.empty((M // W, N), dtype=input.dtype, device=input.device)
    inp = input.contiguous()
    seg_bytes = (M // W) * N * inp.element_size()
    heap = _symm.get_heap()
    st = ctx.state
    _lib.check(_lib.lib().tdist_reduce_scatter(ctypes.byref(heap.ctx), _lib.dtype_code(inp.dtype), inp.data_ptr(),
                                               ctx.scatter_buf.data_ptr(), ctx.scatter_buf.numel() * ctx.scatter_buf.element_size(),
                                               output.data_ptr(), seg_bytes, st.flags_ptr,
                                               st.phase_ptr, st.done_ptr, ctx.num_reduction_sms, _lib.stream_ptr()),
               "reduce_scatter")
    return output


def ring_reduce(input: torch.Tensor, output: torch.Tensor, begin_idx: int, num_splits: int, num_sms: int = -1) -> torch.Tensor:
    """output = sum of the ``num_splits`` equal row-blocks of ``input`` starting after ``begin_idx``."""
    if not input.is_cuda:
        output.copy_(input.view(num_splits, -1, *input.shape[1:]).float().sum(0).to(input.dtype))
        return output
    seg_bytes = input.numel() // num_splits * input.element_size()
    _lib.check(_lib.lib().tdist_slot_reduce(_lib.dtype_code(input.dtype), input.data_ptr(), output.data_ptr(), seg_bytes,
                                            num_splits, begin_idx, 148 if num_sms <= 0 else num_sms, _lib.stream_ptr()),
               "ring_reduce")
    return output
