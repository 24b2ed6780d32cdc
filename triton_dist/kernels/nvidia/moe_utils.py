"""MoE index preparation and top-k reduction (reference ``kernels/nvidia/moe_utils.py:50-405``,
``csrc/lib/moe_utils.cu``). Native kernels: csrc/kernels/moe.cu.

"pair" p = token * topk + k. ``moe_align_sort`` sorts the pairs by expert and pads every expert to ``block_m`` rows so
that each 128-row block of the sorted matrix belongs to exactly one expert (what the grouped GEMM needs)."""
from __future__ import annotations

import dataclasses
from typing import Optional

import torch

from ... import _lib

BLOCK_M = 128


@dataclasses.dataclass
class MoESortResult:
    sorted_pair_ids: torch.Tensor  # [max_rows] int32, pair id per sorted row (num_pairs = padding)
    pair_to_row: torch.Tensor  # [num_pairs] int32, sorted row of every pair
    block_expert: torch.Tensor  # [max_rows / block_m] int32, expert of every row block (-1 = unused)
    meta: torch.Tensor  # [2] int32: rows after padding, number of blocks
    max_rows: int
    num_pairs: int
    topk: int


def max_padded_rows(num_pairs: int, num_experts: int, block_m: int = BLOCK_M) -> int:
    """Upper bound of the padded row count (every expert wastes at most block_m - 1 rows)."""
    rows = num_pairs + min(num_experts, max(num_pairs, 1)) * (block_m - 1)
    return (rows + block_m - 1) // block_m * block_m


def torch_align_sort(topk_ids: torch.Tensor, num_experts: int, block_m: int = BLOCK_M) -> MoESortResult:
    flat = topk_ids.reshape(-1).to(torch.int64)
    num_pairs = flat.numel()
    max_rows = max_padded_rows(num_pairs, num_experts, block_m)
    counts = torch.bincount(flat, minlength=num_experts)
    padded = (counts + block_m - 1) // block_m * block_m
    offs = torch.cumsum(padded, 0) - padded
    order = torch.argsort(flat, stable=True)
    rank_in_e = torch.arange(num_pairs, device=flat.device) - (torch.cumsum(counts, 0) - counts)[flat[order]]
    rows = offs[flat[order]] + rank_in_e
    sorted_ids = torch.full((max_rows,), num_pairs, dtype=torch.int32, device=flat.device)
    sorted_ids[rows] = order.to(torch.int32)
    pair_to_row = torch.empty(num_pairs, dtype=torch.int32, device=flat.device)
    pair_to_row[order] = rows.to(torch.int32)
    block_expert = torch.full((max_rows // block_m,), -1, dtype=torch.int32, device=flat.device)
    total = int(padded.sum())
    be = torch.repeat_interleave(torch.arange(num_experts, device=flat.device), padded // block_m)
    block_expert[:be.numel()] = be.to(torch.int32)
    meta = torch.tensor([total, total // block_m], dtype=torch.int32, device=flat.device)
    return MoESortResult(sorted_ids, pair_to_row, block_expert, meta, max_rows, num_pairs, topk_ids.shape[-1])


def moe_align_sort(topk_ids: torch.Tensor, num_experts: int, block_m: int = BLOCK_M) -> MoESortResult:
    """topk_ids [T, topk] (int32/int64) -> sorted layout (one native single-CTA kernel, no host sync)."""
    if not topk_ids.is_cuda:
        return torch_align_sort(topk_ids, num_experts, block_m)
    ids = topk_ids.reshape(-1).to(torch.int32).contiguous()
    num_pairs = ids.numel()
    max_rows = max_padded_rows(num_pairs, num_experts, block_m)
    dev = ids.device
    res = MoESortResult(torch.empty(max_rows, dtype=torch.int32, device=dev), torch.empty(num_pairs, dtype=torch.int32, device=dev),
                        torch.empty(max_rows // block_m, dtype=torch.int32, device=dev),
                        torch.empty(2, dtype=torch.int32, device=dev), max_rows, num_pairs, topk_ids.shape[-1])
    _lib.check(_lib.lib().tdist_moe_align_sort(ids.data_ptr(), num_pairs, num_experts, block_m, res.sorted_pair_ids.data_ptr(),
                                               res.pair_to_row.data_ptr(), res.block_expert.data_ptr(), res.meta.data_ptr(),
                                               max_rows, _lib.stream_ptr()), "moe_align_sort")
    return res


def moe_gather_rows(x: torch.Tensor, sort: MoESortResult, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [T, H] -> expert-sorted, block-padded rows [max_rows, H] (padding rows are zero)."""
    if out is None:
        out = torch.empty((sort.max_rows, x.shape[1]), dtype=x.dtype, device=x.device)
    if not x.is_cuda:
        out.zero_()
        valid = sort.sorted_pair_ids < sort.num_pairs
        out[valid] = x[(sort.sorted_pair_ids[valid] // sort.topk).long()]
        return out
    _lib.check(_lib.lib().tdist_moe_gather_rows(x.data_ptr(), out.data_ptr(), sort.sorted_pair_ids.data_ptr(), sort.meta.data_ptr(),
                                                sort.num_pairs, sort.topk, x.shape[1] * x.element_size(), sort.max_rows,
                                                _lib.stream_ptr()), "moe_gather_rows")
    return out


def moe_gather_rows_stage(x: torch.Tensor, sort: MoESortResult, block_stage: torch.Tensor, stage: int, out: torch.Tensor,
                          block_m: int = BLOCK_M) -> torch.Tensor:
    """``moe_gather_rows`` restricted to the row blocks b with ``block_stage[b] == stage``; other rows of ``out`` are
    left as they are (the stage-gated AG + grouped GEMM fills ``out`` stage by stage as token shards arrive)."""
    if not x.is_cuda:
        rows = int(sort.meta[0])
        row_stage = torch.repeat_interleave(block_stage[:rows // block_m].long(), block_m)
        ids = sort.sorted_pair_ids[:rows].long()
        sel = row_stage == stage
        pad = sel & (ids >= sort.num_pairs)
        val = sel & (ids < sort.num_pairs)
        out[:rows][pad] = 0
        out[:rows][val] = x[ids[val] // sort.topk]
        return out
    _lib.check(_lib.lib().tdist_moe_gather_rows_stage(x.data_ptr(), out.data_ptr(), sort.sorted_pair_ids.data_ptr(),
                                                      sort.meta.data_ptr(), block_stage.data_ptr(), int(stage), block_m,
                                                      sort.num_pairs, sort.topk, x.shape[1] * x.element_size(), sort.max_rows,
                                                      _lib.stream_ptr()), "moe_gather_rows_stage")
    return out


def reduce_topk(y_sorted: torch.Tensor, sort: MoESortResult, weights: Optional[torch.Tensor] = None,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[t] = sum_k weights[t, k] * y_sorted[row of pair (t, k)] (fp32 accumulate)."""
    tokens = sort.num_pairs // sort.topk
    if out is None:
        out = torch.empty((tokens, y_sorted.shape[1]), dtype=y_sorted.dtype, device=y_sorted.device)
    if not y_sorted.is_cuda or y_sorted.dtype not in (torch.bfloat16, torch.float16, torch.float32):
        rows = y_sorted[sort.pair_to_row.long()].float().view(tokens, sort.topk, -1)
        w = weights.float().view(tokens, sort.topk, 1) if weights is not None else 1.0
        out.copy_((rows * w).sum(1).to(out.dtype))
        return out
    w = weights.float().contiguous() if weights is not None else None
    _lib.check(_lib.lib().tdist_moe_topk_reduce(_lib.dtype_code(y_sorted.dtype), y_sorted.data_ptr(), out.data_ptr(),
                                                sort.pair_to_row.data_ptr(), _lib.ptr(w), tokens, sort.topk,
                                                y_sorted.shape[1] * y_sorted.element_size(), _lib.stream_ptr()), "moe_topk_reduce")
    return out


def grouped_gemm(a_sorted: torch.Tensor, w: torch.Tensor, sort: MoESortResult, out: Optional[torch.Tensor] = None,
                 block_expert: Optional[torch.Tensor] = None, num_sms: int = 0) -> torch.Tensor:
    """y[i] = a_sorted[i] @ w[expert(i)]^T; a_sorted [max_rows, K], w [E, N, K] -> [max_rows, N].
    tcgen05 GEMM engine with a per-128-row-block expert table (padding blocks are skipped on the device).
    ``block_expert`` overrides the table of ``sort`` (entries < 0 = skip the block and leave its rows of ``out`` as they
    are: how the stage-gated AG + MoE computes one arrival stage per launch); ``num_sms`` > 0 limits the grid (leaves SMs to
    a concurrent reduce-scatter kernel)."""
    import ctypes

    E, N, K = w.shape
    fresh = out is None
    if fresh:
        out = torch.empty((sort.max_rows, N), dtype=a_sorted.dtype, device=a_sorted.device)
    if block_expert is None:
        block_expert = sort.block_expert
    if not a_sorted.is_cuda:
        if fresh:
            out.zero_()
        for b in range(sort.max_rows // BLOCK_M):
            e = int(block_expert[b])
            if e >= 0:
                out[b * BLOCK_M:(b + 1) * BLOCK_M] = (a_sorted[b * BLOCK_M:(b + 1) * BLOCK_M].float() @ w[e].float().t()).to(out.dtype)
        return out
    assert a_sorted.is_contiguous() and w.is_contiguous()
    args = _lib.GemmArgs(a=a_sorted.data_ptr(), b=w.data_ptr(), c=out.data_ptr(), M=sort.max_rows, N=N, K=K, lda=K, ldb=K, ldc=N,
                         dtype=_lib.dtype_code(a_sorted.dtype), block_n=0, group_m=1, num_sms=int(num_sms), transpose_out=0,
                         block_expert=block_expert.data_ptr(), num_experts=E, cta_group=1)
    _lib.check(_lib.lib().tdist_gemm(ctypes.byref(args), _lib.stream_ptr()), "grouped_gemm")
    return out
