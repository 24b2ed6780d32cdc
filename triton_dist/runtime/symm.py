"""Symmetric heap: the NVSHMEM replacement.

One VMM allocation per rank, mapped into every rank of the node (see
csrc/runtime/symm_heap.cc), sub-allocated here with a deterministic first-fit
allocator so that the same sequence of ``alloc`` calls yields the same offsets
on all ranks — which is what makes a tensor "symmetric". Reference behaviour
being replaced: ``nvshmem_create_tensor(s)`` / ``nvshmem_free_tensor_sync`` /
``nvshmem_barrier_all_on_stream`` (python/triton_dist/utils.py:114-164).
"""
from __future__ import annotations

import ctypes
import math
import os
import uuid
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from .. import _lib

_ALIGN = 256

_TYPESTR = {
    torch.uint8: "|u1", torch.int8: "|i1", torch.int16: "<i2", torch.int32: "<i4", torch.int64: "<i8",
    torch.float16: "<f2", torch.float32: "<f4", torch.float64: "<f8",
}


class _RawCudaBuffer:
    """Adapter exposing a raw device range through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


def tensor_from_ptr(ptr: int, shape: Sequence[int], dtype: torch.dtype, device: torch.device, owner=None) -> torch.Tensor:
    n = int(math.prod(shape)) * torch.empty((), dtype=dtype).element_size()
    if n == 0:
        return torch.empty(tuple(shape), dtype=dtype, device=device)
    raw = torch.as_tensor(_RawCudaBuffer(ptr, n, owner), device=device)
    return raw.view(dtype).view(tuple(shape))


class SymmetricHeap:
    """Per-process handle of the node-wide symmetric heap."""

    def __init__(self, rank: int, world: int, device: int, nbytes: int, group: Optional[dist.ProcessGroup]):
        self.rank, self.world, self.device_index = rank, world, device
        self.device = torch.device("cuda", device)
        self.group = group
        self._free: List[Tuple[int, int]] = []  # (offset, size) sorted by offset
        self._live: Dict[int, int] = {}  # local ptr -> size
        self._handle = ctypes.c_void_p()
        self.multicast = False
        lib = _lib.lib()
        job = [uuid.uuid4().hex[:16] if rank == 0 else None]
        if world > 1:
            dist.broadcast_object_list(job, src=0, group=group)
        ok = lib.tdist_heap_create(rank, world, device, nbytes, job[0].encode(), ctypes.byref(self._handle)) == 0
        err = "" if ok else lib.tdist_last_error().decode()
        if not self._all_ok(ok):
            # VMM path unavailable on some rank (e.g. fd export not permitted): cudaIpc fallback
            if ok:
                lib.tdist_heap_destroy(self._handle)
            self._handle = ctypes.c_void_p()
            handle = (ctypes.c_ubyte * 64)()
            _lib.check(lib.tdist_heap_create_ipc(rank, world, device, nbytes, ctypes.byref(self._handle), handle),
                       f"symmetric heap (VMM failed with: {err}); cudaIpc fallback")
            if world > 1:
                mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8)
                allh = [torch.empty_like(mine) for _ in range(world)]
                dist.all_gather(allh, mine, group=group)
                blob = (ctypes.c_ubyte * (64 * world))(*[int(x) for t in allh for x in t.tolist()])
                _lib.check(lib.tdist_heap_open_ipc(self._handle, blob), "cudaIpcOpenMemHandle")
        elif os.environ.get("TDIST_DISABLE_NVLS", "0") != "1":
            a_ok = lib.tdist_heap_mc_phase_a(self._handle) == 0
            if self._all_ok(a_ok):
                b_ok = lib.tdist_heap_mc_phase_b(self._handle) == 0
                self.multicast = self._all_ok(b_ok)
        self.nbytes = lib.tdist_heap_bytes(self._handle)
        self.mode = lib.tdist_heap_mode(self._handle)
        self.ctx = _lib.SymmCtx()
        _lib.check(lib.tdist_heap_fill_ctx(self._handle, ctypes.byref(self.ctx)), "fill ctx")
        if not self.multicast:
            self.ctx.mc_base = None
        self._free = [(0, self.nbytes)]
        self.base = [lib.tdist_heap_base(self._handle, p) for p in range(world)]
        self.mc_base = lib.tdist_heap_mc_base(self._handle) if self.multicast else None
        # barrier state lives at the start of the heap
        self._bar_flags = self.alloc((_lib.MAX_RANKS,), torch.int32)
        self._bar_state = self.alloc((2,), torch.int32)
        # The barrier kernel cannot order its own bootstrap: a fast rank's first arrival would land in a slow rank's
        # flag words BEFORE that rank's zero fill above has executed, be wiped by it, and the slow rank would then wait
        # for ever. So every rank finishes zeroing, and all ranks meet on the host, before the first device barrier.
        torch.cuda.synchronize()
        self._host_barrier()
        self.barrier_all()
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ helpers
    def _host_barrier(self) -> None:
        if self.world > 1:
            t = torch.zeros(1, dtype=torch.int32)
            dist.all_reduce(t, group=self.group)

    def _all_ok(self, ok: bool) -> bool:
        if self.world == 1:
            return ok
        t = torch.tensor([1 if ok else 0], dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return bool(t.item())

    # ------------------------------------------------------------------ allocation
    def alloc(self, shape: Sequence[int], dtype: torch.dtype, zero: bool = True) -> torch.Tensor:
        esize = torch.empty((), dtype=dtype).element_size()
        need = max(int(math.prod(shape)) * esize, 1)
        need = (need + _ALIGN - 1) // _ALIGN * _ALIGN
        for i, (off, size) in enumerate(self._free):
            if size >= need:
                if size == need:
                    self._free.pop(i)
                else:
                    self._free[i] = (off + need, size - need)
                ptr = self.base[self.rank] + off
                self._live[ptr] = need
                t = tensor_from_ptr(ptr, shape, dtype, self.device, owner=self)
                if zero:
                    t.zero_()
                return t
        raise MemoryError(f"symmetric heap exhausted: need {need} bytes (heap {self.nbytes}); "
                          "raise TDIST_SYMMETRIC_SIZE")

    def free(self, t: torch.Tensor) -> None:
        ptr = t.data_ptr()
        size = self._live.pop(ptr, None)
        if size is None:
            return
        off = ptr - self.base[self.rank]
        self._free.append((off, size))
        self._free.sort()
        merged: List[Tuple[int, int]] = []
        for o, s in self._free:
            if merged and merged[-1][0] + merged[-1][1] == o:
                merged[-1] = (merged[-1][0], merged[-1][1] + s)
            else:
                merged.append((o, s))
        self._free = merged

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.base[self.rank]
        if not (0 <= off < self.nbytes):
            raise ValueError("tensor does not live in the symmetric heap")
        return off

    def is_symmetric(self, t: torch.Tensor) -> bool:
        return t.is_cuda and 0 <= t.data_ptr() - self.base[self.rank] < self.nbytes

    def peer_view(self, t: torch.Tensor, peer: int) -> torch.Tensor:
        """Tensor aliasing the same symmetric object on ``peer`` (nvshmem get_peer_tensor)."""
        if peer == self.rank:
            return t
        off = self.offset_of(t)
        return tensor_from_ptr(self.base[peer] + off, t.shape, t.dtype, self.device, owner=self)

    def peer_views(self, t: torch.Tensor) -> List[torch.Tensor]:
        return [self.peer_view(t, p) for p in range(self.world)]

    def mc_ptr(self, t: torch.Tensor) -> int:
        if not self.multicast:
            raise RuntimeError("NVLS multicast is not available on this heap")
        return self.mc_base + self.offset_of(t)

    # ------------------------------------------------------------------ sync
    def barrier_all(self, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Stream-ordered, CUDA-graph-capturable barrier across all ranks."""
        st = self._bar_state
        _lib.check(_lib.lib().tdist_barrier_all(ctypes.byref(self.ctx), self._bar_flags.data_ptr(), st.data_ptr(),
                                                st.data_ptr() + 4, _lib.stream_ptr(stream)), "barrier_all")

    def destroy(self) -> None:
        if self._handle:
            torch.cuda.synchronize()
            _lib.lib().tdist_heap_destroy(self._handle)
            self._handle = ctypes.c_void_p()


class HostHeap:
    """CPU stand-in used when no GPU is present (gloo plumbing tests): plain
    tensors, no peer mapping. Ops take their torch.distributed path."""

    multicast = False
    mode = -1

    def __init__(self, rank: int, world: int, group):
        self.rank, self.world, self.group = rank, world, group
        self.device = torch.device("cpu")
        self.ctx = None

    def alloc(self, shape, dtype, zero: bool = True):
        return torch.zeros(tuple(shape), dtype=dtype) if zero else torch.empty(tuple(shape), dtype=dtype)

    def free(self, t):
        return None

    def is_symmetric(self, t):
        return True

    def peer_view(self, t, peer):
        if peer == self.rank:
            return t
        raise RuntimeError("peer views need the CUDA symmetric heap")

    def peer_views(self, t):
        return [self.peer_view(t, self.rank)]

    def barrier_all(self, stream=None):
        if self.world > 1:
            dist.barrier(group=self.group)

    def destroy(self):
        return None


_HEAP = None


def get_heap():
    if _HEAP is None:
        raise RuntimeError("symmetric heap not initialised: call triton_dist.utils.initialize_distributed() first")
    return _HEAP


def heap_initialized() -> bool:
    return _HEAP is not None


def init_heap(rank: int, world: int, device: int, group, nbytes: Optional[int] = None):
    global _HEAP
    if _HEAP is not None:
        return _HEAP
    if nbytes is None:
        nbytes = int(float(os.environ.get("TDIST_SYMMETRIC_SIZE", os.environ.get("NVSHMEM_SYMMETRIC_SIZE", 4 * 2**30))))
    if _lib.use_native():
        _HEAP = SymmetricHeap(rank, world, device, nbytes, group)
    else:
        _HEAP = HostHeap(rank, world, group)
    return _HEAP


def finalize_heap():
    global _HEAP
    if _HEAP is not None:
        _HEAP.destroy()
        _HEAP = None
