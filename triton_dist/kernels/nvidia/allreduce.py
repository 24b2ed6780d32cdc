"""AllReduce over the symmetric heap. API of the reference's
``kernels/nvidia/allreduce.py`` (context :74-135, per-method launchers :685-1098,
``get_auto_allreduce_method`` :1101-1126, ``all_reduce`` :1129-1208).

Kernels: csrc/kernels/collectives.cu. Methods
  one_shot            push to every peer slot + local fp32 sum (latency-optimal without NVLS)
  two_shot            scatter / reduce / broadcast over P2P stores
  one_shot_multimem   NVLS: multimem.ld_reduce of the whole message on every rank  (small messages)
  two_shot_multimem   NVLS: ld_reduce of 1/W + multimem.st of the result          (large messages)
  double_tree, one_shot_tma, two_shot_multimem_st: aliases of the best equivalent above — the tree
  and TMA variants exist in the reference for PCIe boxes / Hopper and have no advantage behind NVSwitch.
"""
from __future__ import annotations

import ctypes
import dataclasses
from typing import Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ...utils import launch_metadata
from ..allreduce import AllReduceMethod
from ..common_ops import CollectiveState

_NATIVE_CODE = {
    AllReduceMethod.OneShot: 1, AllReduceMethod.OneShot_TMA: 1, AllReduceMethod.OneShot_LL: 1,
    AllReduceMethod.TwoShot: 2, AllReduceMethod.DoubleTree: 2,
    AllReduceMethod.OneShot_Multimem: 5, AllReduceMethod.OneShot_Multimem_LL: 5,
    AllReduceMethod.TwoShot_Multimem: 6, AllReduceMethod.TwoShot_Multimem_ST: 6,
}


def workspace_bytes_per_in_byte(world_size: int, method: AllReduceMethod) -> int:
    """Symmetric workspace needed per input byte (both parity halves included)."""
    code = _NATIVE_CODE[method]
    if code == 1:
        return 2 * world_size
    if code == 2:
        return 4  # (scatter + result) x two parities
    return 2


def get_max_chunk_nbytes(workspace_nbytes: int, world_size: int, method: AllReduceMethod) -> int:
    slack = 64 * _lib.MAX_CTAS * world_size if _NATIVE_CODE[method] == 2 else 0  # per-CTA rounding of the scatter area
    n = (workspace_nbytes - slack) // workspace_bytes_per_in_byte(world_size, method)
    return max(n // 16 * 16, 0)


@dataclasses.dataclass
class AllReduceContext:
    workspace_nbytes: int
    rank: int
    world_size: int
    local_world_size: int
    symm_scatter_buf: Optional[torch.Tensor] = None
    state: Optional[CollectiveState] = None
    phase: int = 0

    def __post_init__(self):
        self.local_rank = self.rank % self.local_world_size
        self.node_id = self.rank // self.local_world_size
        self.nnodes = self.world_size // self.local_world_size
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        if self.native:
            self.symm_scatter_buf = heap.alloc((self.workspace_nbytes,), torch.int8, zero=False)
            self.state = CollectiveState()
            self.multimem = bool(heap.multicast)

    def finalize(self):
        if self.native and self.symm_scatter_buf is not None:
            torch.cuda.synchronize()
            _symm.get_heap().free(self.symm_scatter_buf)
            self.state.free()
            self.symm_scatter_buf = None


def create_allreduce_ctx(workspace_nbytes: int, rank: int, world_size: int, local_world_size: int) -> AllReduceContext:
    ctx = AllReduceContext(workspace_nbytes=workspace_nbytes, rank=rank, world_size=world_size,
                           local_world_size=min(local_world_size, world_size))
    _symm.get_heap().barrier_all()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    return ctx


def get_auto_allreduce_method(nbytes: int, multimem: Optional[bool] = None) -> AllReduceMethod:
    """NVLS when bound; one-shot below 512 KB (latency bound), two-shot above (bandwidth bound)."""
    if multimem is None:
        multimem = _symm.heap_initialized() and bool(_symm.get_heap().multicast)
    small = nbytes <= 512 * 1024
    if multimem:
        return AllReduceMethod.OneShot_Multimem if small else AllReduceMethod.TwoShot_Multimem
    return AllReduceMethod.OneShot if small else AllReduceMethod.TwoShot


def _run_straggler(ctx: AllReduceContext, straggler_option) -> None:
    if straggler_option and ctx.rank == straggler_option[0]:
        torch.cuda._sleep(int(straggler_option[1]))


def _native_allreduce(ctx: AllReduceContext, x: torch.Tensor, out: torch.Tensor, method: AllReduceMethod, max_sm: int) -> None:
    heap = _symm.get_heap()
    st = ctx.state
    code = _NATIVE_CODE[method]
    # NVLS methods run in place when the input already lives in the symmetric heap (no staging copy)
    in_place = code >= 5 and heap.is_symmetric(x)
    ws, ws_bytes = (x.data_ptr(), x.numel() * x.element_size()) if in_place else (ctx.symm_scatter_buf.data_ptr(),
                                                                                 ctx.workspace_nbytes)
    _lib.check(_lib.lib().tdist_allreduce(ctypes.byref(heap.ctx), code, _lib.dtype_code(x.dtype), x.data_ptr(),
                                          out.data_ptr(), x.numel() * x.element_size(), ws, ws_bytes, st.flags_ptr,
                                          st.phase_ptr, st.done_ptr, max_sm, int(in_place), _lib.stream_ptr()),
               f"all_reduce({method.name})")


@launch_metadata("all_reduce", lambda x, *_, **__: {"shape": "x".join(map(str, x.shape)), "bytes": x.numel() * x.element_size()})
def all_reduce(x: torch.Tensor, output: Optional[torch.Tensor] = None, method: AllReduceMethod = AllReduceMethod.Unknown,
               ctx: Optional[AllReduceContext] = None, max_sm: int = -1, straggler_option=None) -> torch.Tensor:
    """Sum ``x`` over all ranks into ``output``; large inputs are processed in workspace-sized chunks."""
    if output is None:
        output = torch.empty_like(x)
    if ctx is None or not ctx.native:
        output.copy_(x)
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(output)
        return output
    assert x.is_contiguous() and output.is_contiguous()
    nbytes = x.numel() * x.element_size()
    if method == AllReduceMethod.Unknown:
        method = get_auto_allreduce_method(nbytes, ctx.multimem)
    if _NATIVE_CODE[method] >= 5 and not ctx.multimem:  # NVLS requested but not bound: P2P equivalent
        method = AllReduceMethod.OneShot if _NATIVE_CODE[method] == 5 else AllReduceMethod.TwoShot
    _run_straggler(ctx, straggler_option)
    if max_sm <= 0:
        max_sm = 8 if nbytes <= 256 * 1024 else (32 if nbytes <= 8 * 2**20 else (64 if nbytes <= 32 * 2**20 else 128))
    if nbytes % 16 != 0:  # pad odd sizes through a scratch copy
        pad = (16 - nbytes % 16) // x.element_size()
        xp = torch.cat([x.reshape(-1), x.new_zeros(pad)])
        op = torch.empty_like(xp)
        all_reduce(xp, op, method, ctx, max_sm, None)
        output.reshape(-1).copy_(op[:x.numel()])
        return output
    chunk = get_max_chunk_nbytes(ctx.workspace_nbytes, ctx.world_size, method)
    assert chunk > 0, "all_reduce workspace too small"
    xb, ob = x.reshape(-1).view(torch.uint8), output.reshape(-1).view(torch.uint8)
    for off in range(0, nbytes, chunk):
        n = min(chunk, nbytes - off)
        _native_allreduce(ctx, xb[off:off + n].view(x.dtype), ob[off:off + n].view(x.dtype), method, max_sm)
    return output
