"""Sequence-parallel prefill attention with all-gathered K/V (reference
``kernels/nvidia/sp_ag_attention_intra_node.py``: context :59-99, KV all-gather :105-183, attention :255-425,
``fused_sp_ag_attn_intra_node`` :432-525).

Q, K and V are sharded on the sequence dimension (optionally zig-zag: rank r holds chunks r and 2W-1-r of every
sequence, which balances causal work). Every rank scatters its K/V shard into GLOBAL token order in every peer's
symmetric gather buffer with P2P stores over NVLink (one small kernel per peer), a stream barrier closes the gather,
and a native flash-attention kernel (csrc/kernels/sp_attn.cu) runs the local queries against the gathered K/V using
global positions for the causal mask. Memory per GPU is O(full KV), exactly as in the reference."""
from __future__ import annotations

import dataclasses
import math
import os
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from ... import _lib
from ...runtime import symm as _symm
from ...utils import launch_metadata


@dataclasses.dataclass
class SPAllGatherAttentionContextIntraNode:
    batch_size: int
    q_head: int
    kv_head: int
    max_seqlen_k: int
    max_q_shard_len: int
    head_dim: int
    input_dtype: torch.dtype
    output_dtype: torch.dtype
    rank: int
    world_size: int
    ag_k_buffer: Optional[torch.Tensor] = None
    ag_v_buffer: Optional[torch.Tensor] = None
    attn_output_buffer: Optional[torch.Tensor] = None

    def __post_init__(self):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        shape = (self.batch_size * self.max_seqlen_k, self.kv_head, self.head_dim)
        self.ag_k_buffer = heap.alloc(shape, self.input_dtype, zero=False)
        self.ag_v_buffer = heap.alloc(shape, self.input_dtype, zero=False)
        self.ag_k_buffers = heap.peer_views(self.ag_k_buffer) if self.native else [self.ag_k_buffer]
        self.ag_v_buffers = heap.peer_views(self.ag_v_buffer) if self.native else [self.ag_v_buffer]
        dev = self.ag_k_buffer.device
        self.attn_output_buffer = torch.empty((self.batch_size * self.max_q_shard_len, self.q_head, self.head_dim),
                                              dtype=self.output_dtype, device=dev)
        self.ag_stream = torch.cuda.Stream() if self.native else None
        heap.barrier_all()

    def finalize(self):
        heap = _symm.get_heap()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        heap.free(self.ag_k_buffer)
        heap.free(self.ag_v_buffer)


def create_sp_ag_attention_context_intra_node(batch_size, q_head, kv_head, max_seqlen_k, max_q_shard_len, head_dim, input_dtype,
                                              output_dtype, rank, world_size, device=None):
    return SPAllGatherAttentionContextIntraNode(batch_size, q_head, kv_head, max_seqlen_k, max_q_shard_len, head_dim, input_dtype,
                                                output_dtype, rank, world_size)


def _global_pos(i: torch.Tensor, L: int, rank: int, world: int, zigzag: bool) -> torch.Tensor:
    if not zigzag:
        return rank * (L // world) + i
    h = L // (2 * world)
    return torch.where(i < h, rank * h + i, (2 * world - 1 - rank) * h + (i - h))


def torch_sp_attention(q_shard, k_full, v_full, cu_q, cu_k, rank, world, causal=True, zigzag=True, scale=None):
    """fp32 oracle on already gathered (global-order) K/V."""
    Hq, D = q_shard.shape[1], q_shard.shape[2]
    g = Hq // k_full.shape[1]
    scale = scale or 1.0 / math.sqrt(D)
    out = torch.empty_like(q_shard, dtype=torch.float32)
    for b in range(cu_q.numel() - 1):
        qs, qe, ks, ke = int(cu_q[b]), int(cu_q[b + 1]), int(cu_k[b]), int(cu_k[b + 1])
        L = ke - ks
        q = q_shard[qs:qe].float().transpose(0, 1)  # [Hq, Lq, D]
        k = k_full[ks:ke].float().repeat_interleave(g, dim=1).transpose(0, 1)
        v = v_full[ks:ke].float().repeat_interleave(g, dim=1).transpose(0, 1)
        s = torch.einsum("hqd,hkd->hqk", q, k) * scale
        if causal:
            pos = _global_pos(torch.arange(qe - qs, device=q.device), L, rank, world, zigzag)
            mask = torch.arange(L, device=q.device)[None, :] <= pos[:, None]
            s = s.masked_fill(~mask[None], float("-inf"))
        out[qs:qe] = torch.einsum("hqk,hkd->hqd", torch.softmax(s, -1), v).transpose(0, 1)
    return out


def _gather_kv(ctx, k_shard, v_shard, cu_k, rank, world, zigzag):
    """All-gather K/V into global order on every rank. Native: P2P scatter kernels to every peer + stream barrier."""
    B = cu_k.numel() - 1
    if not ctx.native:
        ks = [torch.zeros_like(k_shard) for _ in range(world)]
        vs = [torch.zeros_like(v_shard) for _ in range(world)]
        if world > 1:
            dist.all_gather(ks, k_shard.contiguous())
            dist.all_gather(vs, v_shard.contiguous())
        else:
            ks, vs = [k_shard], [v_shard]
        for b in range(B):
            s, e = int(cu_k[b]), int(cu_k[b + 1])
            L = e - s
            ls, le = s // world, e // world
            for r in range(world):
                pos = _global_pos(torch.arange(le - ls), L, r, world, zigzag) + s
                ctx.ag_k_buffer[pos] = ks[r][ls:le]
                ctx.ag_v_buffer[pos] = vs[r][ls:le]
        return
    heap = _symm.get_heap()
    lib = _lib.lib()
    row_bytes = k_shard.shape[1] * k_shard.shape[2] * k_shard.element_size()
    max_rows = k_shard.shape[0]
    cur = torch.cuda.current_stream()
    ctx.ag_stream.wait_stream(cur)
    with torch.cuda.stream(ctx.ag_stream):
        for i in range(world):
            peer = (rank + i) % world
            for src, bufs in ((k_shard, ctx.ag_k_buffers), (v_shard, ctx.ag_v_buffers)):
                _lib.check(lib.tdist_sp_kv_scatter(src.data_ptr(), bufs[peer].data_ptr(), cu_k.data_ptr(), B, max_rows, row_bytes, rank,
                                                   world, int(zigzag), _lib.stream_ptr(ctx.ag_stream)), "sp_kv_scatter")
        heap.barrier_all(ctx.ag_stream)  # every rank's rows have landed everywhere
    cur.wait_stream(ctx.ag_stream)


@launch_metadata("flash_attn_varlen", lambda q, k, *_, **__: {"q": "x".join(map(str, q.shape)), "k": "x".join(map(str, k.shape))})
def flash_attn_varlen(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_q: torch.Tensor, cu_k: torch.Tensor, max_q: int,
                      causal: bool = True, out: Optional[torch.Tensor] = None, scale: Optional[float] = None, rank: int = 0,
                      world: int = 1, zigzag: bool = False, k_batch_stride: int = 0) -> torch.Tensor:
    """Varlen GQA flash attention forward on tcgen05 (csrc/kernels/sp_attn_sm100.cu), head_dim 128, bf16.

    q [total_q, Hq, 128], k / v [total_k, Hkv, 128] (token-major; a view with a larger token stride is fine, e.g. a slice of
    a fused qkv tensor), cu_q / cu_k int32 [B+1]. With ``world`` > 1, q holds this rank's share of every sequence and the
    causal mask uses each row's GLOBAL position (plain or zig-zag sharding, as in ``fused_sp_ag_attn_intra_node``); with
    one rank the queries of a sequence are its last rows (prefill: Lq == L; chunked prefill: Lq < L).
    ``k_batch_stride`` > 0: k / v are a dense cache flattened to [B * S, Hkv, 128]; sequence b starts at row b * S and
    cu_k only carries the lengths."""
    total_q, Hq, D = q.shape
    total_k, Hkv = k.shape[0], k.shape[1]
    assert q.is_cuda and q.dtype == torch.bfloat16 and D == 128 and k.dtype == v.dtype == torch.bfloat16
    assert q.stride(2) == 1 and q.stride(1) == D and k.stride(2) == 1 and k.stride(1) == D and v.stride() == k.stride()
    if out is None:
        out = torch.empty((total_q, Hq, D), dtype=q.dtype, device=q.device)
    assert out.is_contiguous()
    _lib.check(_lib.lib().tdist_attn_fwd_tc(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), cu_q.data_ptr(), cu_k.data_ptr(),
                                            cu_q.numel() - 1, max_q, total_q, total_k, Hq, Hkv, D, q.stride(0), k.stride(0), k_batch_stride, rank, world,
                                            int(causal), int(zigzag), float(scale if scale is not None else 1.0 / math.sqrt(D)),
                                            _lib.stream_ptr()), "attn_fwd_tc")
    return out


def fused_sp_ag_attn_intra_node(ctx: SPAllGatherAttentionContextIntraNode, q_shard: torch.Tensor, k_shard: torch.Tensor,
                                v_shard: torch.Tensor, output: Optional[torch.Tensor], cu_seqlens_q: torch.Tensor,
                                cu_seqlens_k: torch.Tensor, max_seqlen_q: int, max_seqlen_k: int, rank: int, world_size: int,
                                is_causal: bool = True, enable_zig_zag: bool = True) -> torch.Tensor:
    """q_shard [total_q_shard, Hq, D]; k_shard/v_shard [total_kv_shard, Hkv, D]; cu_seqlens_q: LOCAL query offsets [B+1]
    (int32); cu_seqlens_k: GLOBAL key offsets [B+1] (every sequence length divisible by W, or 2W with zig-zag).
    Returns output [total_q_shard, Hq, D]."""
    if output is None:
        output = ctx.attn_output_buffer[:q_shard.shape[0]]
    cu_q = cu_seqlens_q.to(torch.int32).contiguous()
    cu_k = cu_seqlens_k.to(torch.int32).contiguous()
    _gather_kv(ctx, k_shard.contiguous(), v_shard.contiguous(), cu_k, rank, world_size, enable_zig_zag)
    total_k = int(cu_k[-1]) if not ctx.native else ctx.ag_k_buffer.shape[0]
    D = q_shard.shape[2]
    scale = 1.0 / math.sqrt(D)
    if not (ctx.native and q_shard.dtype == torch.bfloat16 and D == 128):
        o = torch_sp_attention(q_shard, ctx.ag_k_buffer[:total_k], ctx.ag_v_buffer[:total_k], cu_q, cu_k, rank, world_size,
                               is_causal, enable_zig_zag, scale)
        output.copy_(o.to(output.dtype))
        return output
    B = cu_q.numel() - 1
    if os.environ.get("TDIST_ATTN_LEGACY", "0") != "1":
        return flash_attn_varlen(q_shard, ctx.ag_k_buffer, ctx.ag_v_buffer, cu_q, cu_k, max_seqlen_q, is_causal, out=output,
                                 rank=rank, world=world_size, zigzag=enable_zig_zag)
    _lib.check(_lib.lib().tdist_sp_attn_fwd(q_shard.contiguous().data_ptr(), ctx.ag_k_buffer.data_ptr(), ctx.ag_v_buffer.data_ptr(),
                                            output.data_ptr(), cu_q.data_ptr(), cu_k.data_ptr(), B, max_seqlen_q, q_shard.shape[1],
                                            k_shard.shape[1], D, rank, world_size, int(is_causal), int(enable_zig_zag), scale,
                                            _lib.stream_ptr()), "sp_attn_fwd")
    return output
