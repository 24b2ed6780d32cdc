"""CommOp: point-to-point activation hand-off for pipeline parallelism (reference
``layers/nvidia/p2p.py:43-132`` + ``kernels/nvidia/p2p.py:30-85``).

Double-buffered symmetric activation buffers plus one signal per (buffer, peer). The producer sets the
receiver's signal after its activation is in place; the receiver waits for the signal on its stream and
pulls the rows from the producer's buffer with a P2P copy kernel."""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...kernels.common_ops import CMP_GE, set_signal, wait_eq
from ...runtime import symm as _symm


class CommOp:

    def __init__(self, rank: int, world_size: int, max_tokens: int, hidden: int, dtype: torch.dtype = torch.bfloat16,
                 num_buffers: int = 2, num_sms: int = 4):
        self.rank, self.world_size = rank, world_size
        self.max_tokens, self.hidden, self.dtype = max_tokens, hidden, dtype
        self.num_buffers, self.num_sms = num_buffers, num_sms
        heap = _symm.get_heap()
        self.heap = heap
        self.native = isinstance(heap, _symm.SymmetricHeap)
        self.buffers = heap.alloc((num_buffers, max_tokens, hidden), dtype)
        self.signals = heap.alloc((num_buffers, world_size), torch.int32)
        self.epoch = [0] * num_buffers
        heap.barrier_all()

    def get_buffer(self, buf_id: int, num_tokens: Optional[int] = None) -> torch.Tensor:
        b = self.buffers[buf_id % self.num_buffers]
        return b if num_tokens is None else b[:num_tokens]

    def set_signal(self, peer: int, buf_id: int, value: int, stream=None) -> None:
        """Tell ``peer`` that buffer ``buf_id`` on THIS rank holds data generation ``value``."""
        buf_id %= self.num_buffers
        if not self.native:
            return
        sig = self.heap.peer_view(self.signals, peer)[buf_id, self.rank:self.rank + 1]
        set_signal(sig, value, stream)

    def wait_signal(self, src: int, buf_id: int, value: int, stream=None) -> None:
        buf_id %= self.num_buffers
        if not self.native:
            return
        wait_eq(self.signals[buf_id, src:src + 1], value, stream, cmp=CMP_GE)

    def read(self, src: int, buf_id: int, num_tokens: int, out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
        """Pull ``num_tokens`` rows of buffer ``buf_id`` from rank ``src`` into a local tensor."""
        buf_id %= self.num_buffers
        if out is None:
            out = torch.empty((num_tokens, self.hidden), dtype=self.dtype, device=self.buffers.device)
        if not self.native:
            raise RuntimeError("CommOp.read needs the CUDA symmetric heap")
        remote = self.heap.peer_view(self.buffers, src)[buf_id, :num_tokens]
        nbytes = num_tokens * self.hidden * self.buffers.element_size()
        _lib.check(_lib.lib().tdist_p2p_copy(out.data_ptr(), remote.data_ptr(), nbytes, self.num_sms, _lib.stream_ptr(stream)),
                   "p2p_copy")
        return out

    def finalize(self):
        if self.native:
            torch.cuda.synchronize()
        self.heap.free(self.buffers)
        self.heap.free(self.signals)
