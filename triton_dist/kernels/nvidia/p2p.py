"""P2P copy kernels for pipeline parallelism (reference ``kernels/nvidia/p2p.py:30-85``: ``p2p_copy_kernel`` /
``p2p_copy_remote_to_local_kernel`` built on ``getmem_block``).

Here one native kernel (``tdist_p2p_copy``: 16-byte coherent loads from the peer mapping, ``num_sms`` CTAs) does the
pull; a push is the same kernel with source and destination swapped, because peers are plain mapped addresses."""
from __future__ import annotations

from typing import Optional

import torch

from ... import _lib
from ...runtime import symm as _symm


def _copy(dst: torch.Tensor, src: torch.Tensor, num_sms: int, stream) -> torch.Tensor:
    assert dst.is_contiguous() and src.is_contiguous() and dst.numel() * dst.element_size() == src.numel() * src.element_size()
    nbytes = src.numel() * src.element_size()
    if not src.is_cuda:
        dst.view(-1).copy_(src.view(-1))
        return dst
    _lib.check(_lib.lib().tdist_p2p_copy(dst.data_ptr(), src.data_ptr(), nbytes, num_sms, _lib.stream_ptr(stream)), "p2p_copy")
    return dst


def p2p_copy_remote_to_local(symm_src: torch.Tensor, src_rank: int, out: Optional[torch.Tensor] = None, num_sms: int = 4,
                             stream=None) -> torch.Tensor:
    """Pull the symmetric tensor ``symm_src`` of rank ``src_rank`` into a local tensor."""
    heap = _symm.get_heap()
    remote = heap.peer_view(symm_src, src_rank)
    if out is None:
        out = torch.empty_like(symm_src)
    return _copy(out, remote, num_sms, stream)


def p2p_copy_local_to_remote(local: torch.Tensor, symm_dst: torch.Tensor, dst_rank: int, num_sms: int = 4, stream=None) -> torch.Tensor:
    """Push a local tensor into rank ``dst_rank``'s instance of the symmetric tensor ``symm_dst``."""
    heap = _symm.get_heap()
    return _copy(heap.peer_view(symm_dst, dst_rank), local, num_sms, stream)


# names of the reference's Triton kernels, bound to the launchers
p2p_copy_kernel = p2p_copy_local_to_remote
p2p_copy_remote_to_local_kernel = p2p_copy_remote_to_local
