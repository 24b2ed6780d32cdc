"""Grouped GEMM + top-k reduce + reduce-scatter for tensor-parallel MoE (reference
``kernels/nvidia/moe_reduce_rs.py``: context :60-160, grouped GEMM K46 :166-244, fused reduce/scatter K47 :292-376,
``run_moe_reduce_rs`` :528-660).

Second FFN matrix of every expert is sharded on its input dim ([E, H, I/W]). Two schedules:

* default: tcgen05 grouped GEMM on the expert-sorted activations, then ONE kernel that applies the top-k weights, sums the
  top-k rows of every token and pushes each token segment to its owner rank over NVLink, followed (same kernel) by the
  owner's fp32 sum over ranks.
* ``overlap=True`` (context argument, or ``TDIST_MOE_RS_OVERLAP=1``): the reference's schedule (:569-660): the output
  columns are cut into ``n_chunks`` chunks; the grouped GEMM of chunk c+1 runs on the compute stream (leaving
  ``reduce_ctas`` SMs free) while the reduce + scatter kernel of chunk c runs on a second stream, ordered by one event
  per chunk. The reduce kernel only waits for its PEERS' reduce kernels of the same chunk (never for a later launch on its
  own GPU), so the two streams cannot deadlock. The expert weights are re-laid out once per weight tensor as
  ``n_chunks`` contiguous [E, H / n_chunks, I/W] slabs (cached in the context). Same kernels as the default schedule;
  opt-in because it has not been timed on hardware yet."""
from __future__ import annotations

import ctypes
import dataclasses
import os
from typing import Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ...utils import gemm_metadata, launch_metadata
from ..common_ops import CollectiveState
from .moe_utils import MoESortResult, grouped_gemm, reduce_topk
from .reduce_scatter import torch_reduce_scatter


@dataclasses.dataclass
class MoEReduceRSContext:
    rank: int
    world_size: int
    local_world_size: int
    max_token_num: int
    hidden_dim: int
    num_experts: int
    topk: int
    input_dtype: torch.dtype
    output_dtype: torch.dtype
    n_chunks_max: int = 4
    scatter_buf: Optional[torch.Tensor] = None
    state: Optional[CollectiveState] = None
    overlap: bool = False
    reduce_ctas: int = 16

    def __post_init__(self):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        self.overlap = self.overlap or os.environ.get("TDIST_MOE_RS_OVERLAP", "0") == "1"
        self.rs_stream = torch.cuda.Stream() if (self.overlap and self.native) else None
        self._w_chunks = {}  # (data_ptr, version, n_chunks) -> list of contiguous [E, H / n_chunks, K] slabs
        if self.native:
            self.scatter_buf = heap.alloc((2, self.world_size, self.max_token_num // self.world_size, self.hidden_dim),
                                          self.output_dtype, zero=False)
            self.state = CollectiveState()
            heap.barrier_all()

    def finalize(self):
        if self.native and self.scatter_buf is not None:
            torch.cuda.synchronize()
            _symm.get_heap().free(self.scatter_buf)
            self.state.free()
            self.scatter_buf = None


def create_moe_rs_context(rank, world_size, local_world_size, max_token_num, hidden_dim, num_experts, topk, input_dtype,
                          output_dtype=None, device=None, moe_block_size=128, n_chunks_max=4, overlap=False,
                          **_ignored) -> MoEReduceRSContext:
    return MoEReduceRSContext(rank=rank, world_size=world_size, local_world_size=local_world_size, max_token_num=max_token_num,
                              hidden_dim=hidden_dim, num_experts=num_experts, topk=topk, input_dtype=input_dtype,
                              output_dtype=output_dtype or input_dtype, n_chunks_max=n_chunks_max, overlap=overlap)


def _reduce_rs_launch(ctx: MoEReduceRSContext, y: torch.Tensor, sort: MoESortResult, wts: torch.Tensor, tokens: int,
                      max_ctas: int) -> torch.Tensor:
    out = torch.empty((tokens // ctx.world_size, y.shape[1]), dtype=ctx.output_dtype, device=y.device)
    heap = _symm.get_heap()
    st = ctx.state
    _lib.check(_lib.lib().tdist_moe_reduce_rs(ctypes.byref(heap.ctx), _lib.dtype_code(y.dtype), y.data_ptr(),
                                              ctx.scatter_buf.data_ptr(), ctx.scatter_buf.numel() * ctx.scatter_buf.element_size(),
                                              out.data_ptr(), sort.pair_to_row.data_ptr(),
                                              wts.data_ptr(), tokens, sort.topk, y.shape[1] * y.element_size(), st.flags_ptr,
                                              st.phase_ptr, st.done_ptr, max_ctas, _lib.stream_ptr()), "moe_reduce_rs")
    return out


def _weight_chunks(ctx: MoEReduceRSContext, w: torch.Tensor, n_chunks: int):
    key = (w.data_ptr(), w._version, n_chunks)
    slabs = ctx._w_chunks.get(key)
    if slabs is None:
        ctx._w_chunks.clear()  # one weight tensor per context: do not keep stale copies alive
        slabs = [c.contiguous() for c in w.chunk(n_chunks, dim=1)]
        ctx._w_chunks[key] = slabs
    return slabs


def _pick_chunks(hidden: int, want: int) -> int:
    """Largest chunk count <= want that cuts the hidden dim into equal, 16-byte aligned column blocks of >= 256."""
    for n in range(max(want, 1), 0, -1):
        if hidden % n == 0 and (hidden // n) % 64 == 0 and hidden // n >= 256:
            return n
    return 1


@launch_metadata("moe_reduce_rs", lambda h, w, sort, wts, ctx, *_, **__: gemm_metadata(sort.num_pairs, w.shape[1], w.shape[2], h.element_size(), ctx.world_size,
                                                                                     sort.num_pairs // sort.topk * w.shape[1] * h.element_size()))
def run_moe_reduce_rs(h_sorted: torch.Tensor, w: torch.Tensor, sort: MoESortResult, topk_weights: torch.Tensor,
                      ctx: MoEReduceRSContext, n_chunks: int = 4) -> torch.Tensor:
    """h_sorted [max_rows, I/W] (expert-sorted), w [E, H, I/W], topk_weights [M, topk] for ALL tokens.
    Returns this rank's token segment [M/W, H]."""
    tokens = sort.num_pairs // sort.topk
    W = ctx.world_size
    wts = topk_weights.float().contiguous()
    chunks = _pick_chunks(w.shape[1], min(n_chunks, ctx.n_chunks_max)) if ctx.overlap else 1
    if chunks > 1:
        return _run_moe_reduce_rs_chunked(h_sorted, w, sort, wts, ctx, tokens, chunks)
    y = grouped_gemm(h_sorted, w, sort)  # [max_rows, H] partial over the FFN shard
    if not ctx.native:
        full = reduce_topk(y, sort, wts)
        return torch_reduce_scatter(full) if W > 1 else full
    return _reduce_rs_launch(ctx, y, sort, wts, tokens, 64)


def _run_moe_reduce_rs_chunked(h_sorted, w, sort, wts, ctx: MoEReduceRSContext, tokens: int, chunks: int) -> torch.Tensor:
    """Column-chunk pipeline: GEMM of chunk c+1 on the compute stream, reduce + scatter of chunk c on ``ctx.rs_stream``."""
    W = ctx.world_size
    slabs = _weight_chunks(ctx, w, chunks)
    outs = []
    if not ctx.native:  # CPU plumbing: same chunking, no streams
        for wc in slabs:
            full = reduce_topk(grouped_gemm(h_sorted, wc, sort), sort, wts)
            outs.append(torch_reduce_scatter(full) if W > 1 else full)
        return torch.cat(outs, dim=1)
    main = torch.cuda.current_stream()
    gemm_sms = max(torch.cuda.get_device_properties(h_sorted.device).multi_processor_count - ctx.reduce_ctas, 1)
    ctx.rs_stream.wait_stream(main)  # the previous op on the compute stream may still be using the scatter workspace
    for wc in slabs:
        y = grouped_gemm(h_sorted, wc, sort, num_sms=gemm_sms)
        done = torch.cuda.Event()
        done.record(main)
        with torch.cuda.stream(ctx.rs_stream):
            ctx.rs_stream.wait_event(done)
            outs.append(_reduce_rs_launch(ctx, y, sort, wts, tokens, ctx.reduce_ctas))
        y.record_stream(ctx.rs_stream)  # allocated on the compute stream, read on the other one
    main.wait_stream(ctx.rs_stream)
    out = torch.cat(outs, dim=1)
    for t in outs:
        t.record_stream(main)  # allocated on the reduce stream, read by the concatenation on the compute stream
    return out
