"""Token-level expert-parallel dispatch / combine (reference ``kernels/nvidia/ep_a2a.py``: dispatch K38 :36-148,
combine K39 :151-240, splits/offsets K40 :243-306, bincount K41 :309-334).

Single NVSwitch node, so only the intra-node stage exists: (1) per-expert counts are all-gathered with direct stores,
(2) every rank derives the exact receive layout (rows ordered by local expert, then source rank) and stores each
(token, k) row at its final position in the destination rank's buffer — optionally quantised to fp8 e4m3 with one fp32
scale per 128 values, (3) combine pulls the top-k expert outputs of every token from the peers and sums them in fp32.
No host polling: the receive buffer has a fixed capacity and the received count stays on the device
(``num_recv``); ``EPAll2AllLayer`` offers an explicit ``.item()`` for callers that want an exact-size tensor."""
from __future__ import annotations

import ctypes
import dataclasses
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ...utils import all_to_all_compat
from ...utils import launch_metadata
from ..common_ops import CollectiveState


@dataclasses.dataclass
class EPContext:
    rank: int
    world_size: int
    max_tokens: int
    hidden: int
    topk: int
    num_experts: int
    dtype: torch.dtype
    fp8: bool = False
    capacity: int = 0
    num_sm: int = 32

    def __post_init__(self):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        self.experts_per_rank = self.num_experts // self.world_size
        if self.capacity <= 0:  # worst case: every pair of every rank lands here
            self.capacity = self.max_tokens * self.topk * self.world_size
        W, E = self.world_size, self.num_experts
        self.full_counts = heap.alloc((W, E), torch.int32)
        payload_dtype = torch.uint8 if self.fp8 else self.dtype
        self.recv_buf = heap.alloc((self.capacity, self.hidden), payload_dtype, zero=False)
        self.recv_scale = heap.alloc((self.capacity, max(self.hidden // 128, 1)), torch.float32, zero=False)
        self.recv_src_pair = heap.alloc((self.capacity,), torch.int32, zero=False)
        self.expert_out = heap.alloc((self.capacity, self.hidden), self.dtype, zero=False)  # combine source
        dev = self.recv_buf.device
        self.scatter_idx = torch.empty((self.max_tokens * self.topk,), dtype=torch.int32, device=dev)
        self.cursors = torch.zeros((E,), dtype=torch.int32, device=dev)
        self.recv_expert_counts = torch.zeros((self.experts_per_rank,), dtype=torch.int32, device=dev)
        self.num_recv = torch.zeros((1,), dtype=torch.int32, device=dev)
        self.grid_counter = torch.zeros((4,), dtype=torch.int32, device=dev)
        self.state = CollectiveState(4 * _lib.MAX_RANKS) if self.native else None
        heap.barrier_all()

    def finalize(self):
        heap = _symm.get_heap()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        for t in (self.full_counts, self.recv_buf, self.recv_scale, self.recv_src_pair, self.expert_out):
            heap.free(t)
        if self.state is not None:
            self.state.free()


def _torch_dispatch(ctx: EPContext, x: torch.Tensor, ids: torch.Tensor):
    """gloo reference: same receive layout (local expert major, source rank minor; arrival order inside)."""
    W, E, epr, r = ctx.world_size, ctx.num_experts, ctx.experts_per_rank, ctx.rank
    T, topk = ids.shape
    flat = ids.reshape(-1).long()
    counts = torch.bincount(flat, minlength=E).int()
    allc = [torch.zeros_like(counts) for _ in range(W)]
    if W > 1:
        dist.all_gather(allc, counts)
    else:
        allc = [counts]
    full = torch.stack(allc)  # [W, E]
    ctx.full_counts.copy_(full)
    order = torch.argsort(flat, stable=True)
    send_rows = x[(order // topk)]
    dst = flat[order] // epr
    ins = [send_rows[dst == d] for d in range(W)]
    in_meta = [torch.stack([flat[order][dst == d], order[dst == d]], 1) for d in range(W)]
    rc = [int(full[s, r * epr:(r + 1) * epr].sum()) for s in range(W)]
    outs = [x.new_empty((n, ctx.hidden)) for n in rc]
    out_meta = [torch.empty((n, 2), dtype=torch.long) for n in rc]
    if W > 1:
        all_to_all_compat(outs, ins)
        all_to_all_compat(out_meta, in_meta)
    else:
        outs, out_meta = ins, in_meta
    # place: rows ordered by (local expert, src)
    base = torch.zeros((epr, W), dtype=torch.long)
    acc = 0
    for e in range(epr):
        for s in range(W):
            base[e, s] = acc
            acc += int(full[s, r * epr + e])
    slots_per_src = []
    for s in range(W):
        e_loc = out_meta[s][:, 0] - r * epr
        within = torch.zeros_like(e_loc)
        for e in range(epr):
            m = e_loc == e
            within[m] = torch.arange(int(m.sum()))
        slot = base[e_loc, s] + within
        ctx.recv_buf[slot] = outs[s].to(ctx.recv_buf.dtype)
        ctx.recv_src_pair[slot] = (s * T * topk + out_meta[s][:, 1]).int()
        slots_per_src.append(slot)
    # tell every source where its pairs went
    back = [torch.empty((len(in_meta[d]),), dtype=torch.long) for d in range(W)]
    if W > 1:
        all_to_all_compat(back, slots_per_src)
    else:
        back = slots_per_src
    for d in range(W):
        ctx.scatter_idx[in_meta[d][:, 1]] = back[d].int()
    ctx.recv_expert_counts.copy_(full[:, r * epr:(r + 1) * epr].sum(0).int())
    ctx.num_recv.fill_(acc)


@launch_metadata("ep_dispatch", lambda ctx, x, ids, *_, **__: {"tokens": x.shape[0], "hidden": x.shape[1], "topk": ids.shape[-1], "W": ctx.world_size})
def ep_dispatch(ctx: EPContext, x: torch.Tensor, topk_ids: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """x [T, hidden], topk_ids [T, topk] (global expert ids). Returns (recv_buf [capacity, hidden] (first num_recv rows
    valid, ordered by local expert), recv_expert_counts [experts_per_rank], num_recv [1]) — all on the device."""
    T, topk = topk_ids.shape
    assert T <= ctx.max_tokens and topk == ctx.topk and x.shape[1] == ctx.hidden
    ids = topk_ids.to(torch.int32).contiguous()
    ctx._last_ids, ctx._last_tokens = ids, T
    if not ctx.native:
        _torch_dispatch(ctx, x, ids)
        return ctx.recv_buf, ctx.recv_expert_counts, ctx.num_recv
    heap = _symm.get_heap()
    st = ctx.state
    lib = _lib.lib()
    _lib.check(lib.tdist_ep_counts(ctypes.byref(heap.ctx), ids.data_ptr(), T * topk, ctx.num_experts, ctx.full_counts.data_ptr(),
                                   st.flags_ptr, st.phase_ptr, st.done_ptr, _lib.stream_ptr()), "ep_counts")
    xc = x.contiguous()
    _lib.check(lib.tdist_ep_dispatch(ctypes.byref(heap.ctx), _lib.dtype_code(x.dtype), xc.data_ptr(), ids.data_ptr(),
                                     ctx.full_counts.data_ptr(), ctx.recv_buf.data_ptr(), ctx.recv_scale.data_ptr(),
                                     ctx.recv_src_pair.data_ptr(), ctx.scatter_idx.data_ptr(), ctx.cursors.data_ptr(),
                                     ctx.recv_expert_counts.data_ptr(), ctx.num_recv.data_ptr(), T, topk, ctx.hidden,
                                     ctx.num_experts, int(ctx.fp8), st.flags_ptr + 16 * 4, ctx.grid_counter.data_ptr(),
                                     st.phase_ptr, st.done_ptr, ctx.num_sm, _lib.stream_ptr()), "ep_dispatch")
    return ctx.recv_buf, ctx.recv_expert_counts, ctx.num_recv


@launch_metadata("ep_combine", lambda ctx, y, *_, **__: {"rows": y.shape[0], "hidden": y.shape[1], "W": ctx.world_size})
def ep_combine(ctx: EPContext, expert_out: torch.Tensor, topk_weights: Optional[torch.Tensor] = None) -> torch.Tensor:
    """expert_out [>= num_recv, hidden] in dispatch row order -> [T, hidden] = sum_k w[t,k] * expert_out(pair (t,k))."""
    T, ids = ctx._last_tokens, ctx._last_ids
    n = min(expert_out.shape[0], ctx.capacity)
    if expert_out.data_ptr() != ctx.expert_out.data_ptr():
        ctx.expert_out[:n].copy_(expert_out[:n])
    out = torch.empty((T, ctx.hidden), dtype=ctx.dtype, device=expert_out.device)
    w = topk_weights.float().contiguous() if topk_weights is not None else None
    if not ctx.native:
        W, epr = ctx.world_size, ctx.experts_per_rank
        gathered = [torch.zeros_like(ctx.expert_out) for _ in range(W)]
        if W > 1:
            dist.all_gather(gathered, ctx.expert_out)
        else:
            gathered = [ctx.expert_out]
        allbuf = torch.stack(gathered)  # [W, cap, H]
        flat = ids.reshape(-1).long()
        rows = allbuf[flat // epr, ctx.scatter_idx[:flat.numel()].long()].float().view(T, ctx.topk, -1)
        ww = w.view(T, ctx.topk, 1) if w is not None else 1.0
        out.copy_((rows * ww).sum(1).to(out.dtype))
        return out
    heap = _symm.get_heap()
    st = ctx.state
    _lib.check(_lib.lib().tdist_ep_combine(ctypes.byref(heap.ctx), _lib.dtype_code(ctx.dtype), ctx.expert_out.data_ptr(),
                                           ids.data_ptr(), ctx.scatter_idx.data_ptr(), _lib.ptr(w), out.data_ptr(), T, ctx.topk,
                                           ctx.hidden, ctx.num_experts, st.flags_ptr + 32 * 4,
                                           ctx.grid_counter.data_ptr() + 8, st.phase_ptr, st.done_ptr, 64, _lib.stream_ptr()),
               "ep_combine")
    return out
