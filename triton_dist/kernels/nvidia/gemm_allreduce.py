"""GEMM + AllReduce (tensor-parallel row linear, replicated output).

API of the reference's ``kernels/nvidia/gemm_allreduce.py`` (contexts :46-120,
``low_latency_gemm_allreduce_op`` :509-543, ``gemm_allreduce_op`` :546-575).

B200 implementation: ONE launch of the tcgen05 GEMM engine in mode AR for both
entry points. GEMM CTAs write bf16 tiles into a symmetric buffer and release a
per-tile flag; a few comm CTAs wait until every rank has published a tile and
reduce it across ranks with NVLS ``multimem.ld_reduce`` (P2P loads when
multicast is unavailable). "One-shot" (every rank reduces every tile; decode
sizes) and "two-shot" (rank r reduces tiles r, r+W, ... and multimem.st's them
to everybody; prefill sizes) are selected by output size. The tile buffer is
double buffered by call parity — the role of the reference's ``num_phases=2``
context pair — so there is no trailing barrier and the op is graph-capturable.
"""
from __future__ import annotations

import ctypes
import os
import dataclasses
from typing import List, Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ..common_ops import CollectiveState
from .allgather_gemm import _check_operands, _gemm_args

_MIN_BN = 32


@dataclasses.dataclass
class GemmARContext:
    rank: int
    world_size: int
    max_M: int
    N: int
    dtype: torch.dtype
    NUM_COMM_SMS: int = 8
    ar_stream: Optional["torch.cuda.Stream"] = None
    symm_gemm_out_buf: Optional[torch.Tensor] = None  # [2, max_M, N]
    symm_ar_out_buf: Optional[torch.Tensor] = None  # [max_M, N]
    state: Optional[CollectiveState] = None
    flags2: Optional[torch.Tensor] = None

    def __post_init__(self):
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        if self.native:
            self.symm_gemm_out_buf = heap.alloc((2, self.max_M, self.N), self.dtype, zero=False)
            self.symm_ar_out_buf = heap.alloc((self.max_M, self.N), self.dtype, zero=False)
            n_tiles = ((self.max_M + 127) // 128) * ((self.N + _MIN_BN - 1) // _MIN_BN)
            self.state = CollectiveState(self.world_size * n_tiles + 64)  # one flag per (source rank, tile)
            self.flags2 = heap.alloc((n_tiles + 64,), torch.int32)
            self.use_mc = bool(heap.multicast)
            self.gemv_state = CollectiveState(16 * _lib.MAX_RANKS)  # decode path: GEMV + in-place NVLS all-reduce

    def get_gemm_out_buf(self, input, weight):
        M, N = input.shape[0], weight.shape[0]
        return self.symm_gemm_out_buf[0, :M, :N]

    def finalize(self):
        if self.native and self.symm_gemm_out_buf is not None:
            torch.cuda.synchronize()
            heap = _symm.get_heap()
            heap.free(self.symm_gemm_out_buf)
            heap.free(self.symm_ar_out_buf)
            heap.free(self.flags2)
            self.state.free()
            self.gemv_state.free()
            self.symm_gemm_out_buf = None


@dataclasses.dataclass
class LLGemmARContext:
    """Kept for API parity: the reference alternates ``num_phases`` full contexts;
    here one context double-buffers internally, so all phases alias it."""

    ctxs: List[GemmARContext]
    num_phases: int = 2
    phase: int = 0

    def update_phase(self):
        self.phase = (self.phase + 1) % self.num_phases

    def __getattr__(self, name):
        return getattr(self.ctxs[0], name)

    def finalize(self):
        self.ctxs[0].finalize()


def create_gemm_ar_context(ar_stream, rank, world_size, local_world_size, max_M, N, dtype, MIN_BLOCK_SIZE_M=16,
                           MIN_BLOCK_SIZE_N=16, NUM_COMM_SMS=8) -> GemmARContext:
    assert local_world_size == world_size, "single NVSwitch node only"
    ctx = GemmARContext(rank=rank, world_size=world_size, max_M=max_M, N=N, dtype=dtype,
                        NUM_COMM_SMS=min(NUM_COMM_SMS, 32), ar_stream=ar_stream)
    _symm.get_heap().barrier_all()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    return ctx


def create_ll_gemm_ar_context(rank, world_size, local_world_size, max_M, N, dtype, MIN_BLOCK_SIZE_M=16, MIN_BLOCK_SIZE_N=16,
                              NUM_COMM_SMS=8, num_phases=2) -> LLGemmARContext:
    ctx = create_gemm_ar_context(None, rank, world_size, local_world_size, max_M, N, dtype, NUM_COMM_SMS=NUM_COMM_SMS)
    return LLGemmARContext(ctxs=[ctx] * num_phases, num_phases=num_phases, phase=0)


def torch_gemm_ar(a: torch.Tensor, b: torch.Tensor, group=None) -> torch.Tensor:
    out = torch.matmul(a, b.t())
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out, group=group)
    return out


def _launch(ctx: GemmARContext, a: torch.Tensor, b: torch.Tensor, one_shot: bool, copy_to_local: bool, block_n: int = 0):
    _check_operands(a, b)
    M, K = a.shape
    N = b.shape[0]
    assert M <= ctx.max_M and N == ctx.N, f"context built for max_M={ctx.max_M}, N={ctx.N}"
    if not ctx.native:
        return torch_gemm_ar(a, b)
    heap = _symm.get_heap()
    if ctx.world_size == 1:  # nothing to reduce: the plain linear (GEMV for decode sizes, tcgen05 GEMM otherwise)
        from .elementwise import linear

        return linear(a, b)
    if M <= 16 and one_shot and ctx.use_mc and K % 8 == 0 and os.environ.get("TDIST_GEMM_AR_DECODE", "fused") == "gemv":
        # decode: the tensor core would idle at <= 16 rows. Weight-streaming GEMV straight into the symmetric buffer,
        # then the NVLS one-shot all-reduce in place (multimem.ld_reduce; no staging copy).
        from .elementwise import gemv

        sym = ctx.symm_gemm_out_buf.view(-1)[:M * N].view(M, N)
        gemv(a, b, out=sym)
        out = torch.empty((M, N), dtype=a.dtype, device=a.device)
        st = ctx.gemv_state
        _lib.check(_lib.lib().tdist_allreduce(ctypes.byref(heap.ctx), 5, _lib.dtype_code(a.dtype), sym.data_ptr(), out.data_ptr(),
                                              M * N * a.element_size(), sym.data_ptr(), M * N * a.element_size(), st.flags_ptr,
                                              st.phase_ptr, st.done_ptr, 8, 1, _lib.stream_ptr()), "gemv_allreduce")
        return out
    if one_shot:
        out = torch.empty((M, N), dtype=a.dtype, device=a.device)
    else:
        out = ctx.symm_ar_out_buf[:M]
    # tiles are laid out with row stride N in the symmetric buffer; its two parity halves are max_M * N apart
    sym_c = ctx.symm_gemm_out_buf
    if block_n == 0 and M <= 16:
        block_n = int(os.environ.get("TDIST_GEMM_AR_DECODE_BN", 0))  # sweep knob for the decode-size tile width
    args = _gemm_args(a, b, sym_c[0], M, N, K, block_n=block_n, group_m=8)
    st = ctx.state
    f = _lib.FusedArgs(ctx=heap.ctx, mode=3, ws=sym_c.data_ptr(), out=out.data_ptr(), flags=st.flags_ptr,
                       flags2=ctx.flags2.data_ptr(), counters=None, phase=st.phase_ptr, done=st.done_ptr,
                       num_comm_ctas=ctx.NUM_COMM_SMS, ar_one_shot=int(one_shot), use_mc=int(ctx.use_mc), variant=0,
                       ws_bytes=sym_c.numel() * sym_c.element_size(), flags_words=st.flags.numel())
    _lib.check(_lib.lib().tdist_gemm_fused(ctypes.byref(args), ctypes.byref(f), _lib.stream_ptr()), "gemm_ar")
    if not one_shot and copy_to_local:
        return out.clone()
    return out


def low_latency_gemm_allreduce_op(ctx, a, b, gemm_config=None, copy_to_local=True, USE_MULTIMEM_ST=True):
    """Decode-sized GEMM+AR: every rank reduces every tile as soon as all ranks published it."""
    if isinstance(ctx, LLGemmARContext):
        ctx.update_phase()
        ctx = ctx.ctxs[0]
    return _launch(ctx, a, b, one_shot=True, copy_to_local=copy_to_local)


def gemm_allreduce_op(ctx, a, b, gemm_config=None, copy_to_local=True, USE_MULTIMEM_ST=True):
    """Throughput GEMM+AR: two-shot (each tile reduced by one rank and multicast) for large outputs."""
    if isinstance(ctx, LLGemmARContext):
        ctx = ctx.ctxs[0]
    M, N = a.shape[0], b.shape[0]
    one_shot = ctx.world_size == 1 or M * N * a.element_size() <= 512 * 1024
    return _launch(ctx, a, b, one_shot=one_shot, copy_to_local=copy_to_local)
