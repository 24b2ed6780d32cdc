"""Low-latency expert-parallel all-to-all (reference ``kernels/nvidia/low_latency_all_to_all.py``: kernel :35-118,
context :124-196, ``fast_all_to_all`` :198-257, ``all_to_all_post_process`` :260-279).

Tokens are pre-sorted by (global) expert; rank p receives the rows of its experts from every source into a fixed slot
``[src][max_m]`` of a receive buffer that is double-buffered by call parity, together with the per-expert splits and
optional per-token fp32 scales (fp8 payloads). One kernel, several CTAs per peer, stores over NVLink, one
release/acquire flag exchange; graph-capturable."""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ...utils import all_to_all_compat
from ..common_ops import CollectiveState


class AllToAllContext:

    def __init__(self, max_m: int, hidden: int, rank: int, num_tot_experts: int, WORLD_SIZE: int, experts_per_rank: int,
                 dtype=torch.bfloat16, scale_dtype=torch.float, ctas_per_peer: int = 4):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        self.max_m, self.hidden, self.dtype, self.scale_dtype = max_m, hidden, dtype, scale_dtype
        self.num_tot_experts, self.experts_per_rank = num_tot_experts, experts_per_rank
        self.WORLD_SIZE, self.rank = WORLD_SIZE, rank
        self.ctas_per_peer = ctas_per_peer
        self.call_count = 1
        self.recv_buf = heap.alloc((2, WORLD_SIZE * max_m, hidden), dtype, zero=False)
        self.scale_recv_buf = heap.alloc((2, WORLD_SIZE * max_m), torch.float32, zero=False)
        self.split_recv_buf = heap.alloc((2, num_tot_experts), torch.int32)
        self.state = CollectiveState(2 * _lib.MAX_RANKS) if self.native else None
        heap.barrier_all()

    def finalize(self):
        heap = _symm.get_heap()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        for t in (self.recv_buf, self.scale_recv_buf, self.split_recv_buf):
            heap.free(t)
        if self.state is not None:
            self.state.free()


def create_all_to_all_context(max_m: int, hidden: int, rank: int, num_tot_experts: int, WORLD_SIZE: int, experts_per_rank: int,
                              dtype=torch.bfloat16, scale_dtype=torch.float) -> AllToAllContext:
    return AllToAllContext(max_m, hidden, rank, num_tot_experts, WORLD_SIZE, experts_per_rank, dtype, scale_dtype)


def _torch_all_to_all(ctx: AllToAllContext, send: torch.Tensor, cumsum: torch.Tensor, scale: Optional[torch.Tensor], parity: int):
    W, epr = ctx.WORLD_SIZE, ctx.experts_per_rank
    splits = (cumsum[1:] - cumsum[:-1]).to(torch.int32)
    per_rank = splits.view(W, epr).sum(1)
    all_counts = [torch.zeros_like(per_rank) for _ in range(W)]
    all_splits = [torch.zeros_like(splits) for _ in range(W)]
    if W > 1:
        dist.all_gather(all_counts, per_rank)
        dist.all_gather(all_splits, splits)
    else:
        all_counts, all_splits = [per_rank], [splits]
    recv_counts = [int(c[ctx.rank]) for c in all_counts]
    outs = [send.new_empty((n, ctx.hidden)) for n in recv_counts]
    ins = list(send[:int(cumsum[-1])].split([int(x) for x in per_rank]))
    if W > 1:
        all_to_all_compat(outs, ins)
    else:
        outs = ins
    for s in range(W):
        ctx.recv_buf[parity, s * ctx.max_m:s * ctx.max_m + recv_counts[s]] = outs[s]
        ctx.split_recv_buf[parity, s * epr:(s + 1) * epr] = all_splits[s][ctx.rank * epr:(ctx.rank + 1) * epr]
    if scale is not None:
        souts = [scale.new_empty((n,)) for n in recv_counts]
        sins = list(scale[:int(cumsum[-1])].split([int(x) for x in per_rank]))
        if W > 1:
            all_to_all_compat(souts, sins)
        else:
            souts = sins
        for s in range(W):
            ctx.scale_recv_buf[parity, s * ctx.max_m:s * ctx.max_m + recv_counts[s]] = souts[s].float()


def fast_all_to_all(ctx: AllToAllContext, send_tensor: torch.Tensor, send_split_cumsum: torch.Tensor,
                    send_scale: Optional[torch.Tensor]) -> List[Optional[torch.Tensor]]:
    """send_tensor [T, hidden] sorted by expert; send_split_cumsum [num_tot_experts + 1] int32.
    Returns [recv_splits [W * epr], recv_data [W * max_m, hidden], recv_scale [W * max_m] | None] (views of the
    parity buffer of this call: valid until the call after next)."""
    assert send_tensor.shape[0] <= ctx.max_m
    parity = ctx.call_count & 1
    if not ctx.native:
        _torch_all_to_all(ctx, send_tensor, send_split_cumsum, send_scale, parity)
    else:
        heap = _symm.get_heap()
        st = ctx.state
        cs = send_split_cumsum.to(torch.int32).contiguous()
        sc = send_scale.float().contiguous() if send_scale is not None else None
        send = send_tensor.contiguous()
        _lib.check(_lib.lib().tdist_a2a(ctypes.byref(heap.ctx), send.data_ptr(), cs.data_ptr(), _lib.ptr(sc), ctx.recv_buf.data_ptr(),
                                        ctx.split_recv_buf.data_ptr(), ctx.scale_recv_buf.data_ptr(), st.flags_ptr, st.phase_ptr,
                                        st.done_ptr, ctx.max_m, ctx.hidden * send.element_size(), ctx.experts_per_rank,
                                        ctx.ctas_per_peer, _lib.stream_ptr()), "fast_all_to_all")
    ctx.call_count += 1
    return [ctx.split_recv_buf[parity], ctx.recv_buf[parity], ctx.scale_recv_buf[parity] if send_scale is not None else None]


def all_to_all_post_process(ctx: AllToAllContext, input_splits: torch.Tensor, recv_buffer: torch.Tensor,
                            scale_buffer: Optional[torch.Tensor] = None):
    """Compact the fixed-slot receive buffer into one contiguous [sum(counts), hidden] tensor (+ scales)."""
    W = ctx.WORLD_SIZE
    counts = input_splits.reshape(W, -1).sum(dim=1).tolist()
    data = torch.cat([recv_buffer[i * ctx.max_m:i * ctx.max_m + int(n)] for i, n in enumerate(counts)])
    scale = None
    if scale_buffer is not None:
        scale = torch.cat([scale_buffer[i * ctx.max_m:i * ctx.max_m + int(n)] for i, n in enumerate(counts)])
    return data, scale
