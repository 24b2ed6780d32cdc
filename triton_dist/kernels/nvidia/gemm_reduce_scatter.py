"""GEMM + ReduceScatter (tensor-parallel row linear with sequence-sharded output).

API of the reference's ``kernels/nvidia/gemm_reduce_scatter.py`` (context
:40-82, ``gemm_rs`` :508-583).

B200 implementation: ONE launch of the tcgen05 GEMM engine in mode RS.

* NVLS (default when the heap has a multicast mapping): every rank walks the
  output tiles in the same owner-interleaved order; the epilogue stores each
  partial tile (TMEM -> bf16 -> 128-byte rows) into a LOCAL symmetric buffer
  and flags the owner; the two idle warps of every GEMM CTA on the owner sum
  finished tiles across ranks in the switch (``multimem.ld_reduce``, fp32
  accumulation) while later tiles are still being computed. Nothing is
  scattered, no SM is reserved, there is no reduction tail.
* point-to-point: tiles are produced owner-by-owner starting at ``rank+1``; the
  epilogue stores each partial tile into slot ``rank`` of the owner's scatter
  buffer over NVLink, and after its tiles every CTA joins the flag-gated fp32
  reduction of this rank's own segment.

``fuse_scatter`` is always on; the flag is accepted for API compatibility.
Buffers are double buffered by call parity (fixed distance between the halves),
so no inter-call barrier is needed.
"""
from __future__ import annotations

import ctypes
import dataclasses
from typing import Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ...utils import gemm_metadata, launch_metadata
from .allgather_gemm import _check_operands, _gemm_args
from .reduce_scatter import ReduceScatter2DContext, create_reduce_scater_2d_ctx, torch_reduce_scatter


@dataclasses.dataclass
class GEMMReduceScatterTensorParallelContext:
    rs_ctx: ReduceScatter2DContext
    output_dtype: torch.dtype
    rs_stream: Optional["torch.cuda.Stream"] = None
    num_gemm_sms: int = 0
    BLOCK_M: int = 128
    BLOCK_N: int = 0
    BLOCK_K: int = 64
    GROUP_M: int = 8
    stages: int = 4
    counters: Optional[torch.Tensor] = None
    variant: int = 0  # 0 auto, 1 point-to-point scatter, 2 NVLS pull-reduce
    cta_group: int = 0

    def __post_init__(self):
        if self.rs_ctx.native:
            blocks = self.rs_ctx.world_size * ((self.rs_ctx.max_M // self.rs_ctx.world_size + 127) // 128)
            self.counters = torch.zeros((blocks + 64,), dtype=torch.int32, device="cuda")
            self.num_gemm_sms = torch.cuda.get_device_properties("cuda").multi_processor_count

    def finalize(self):
        self.rs_ctx.finalize()


def create_gemm_rs_context(max_M, N, rank, world_size, local_world_size, output_dtype: torch.dtype,
                           rs_stream: Optional["torch.cuda.Stream"] = None, BLOCK_M=128, BLOCK_N=0, BLOCK_K=64, GROUP_M=8,
                           stages=4) -> GEMMReduceScatterTensorParallelContext:
    rs_ctx = create_reduce_scater_2d_ctx(max_M, N, rank, world_size, local_world_size, output_dtype, overlap_with_gemm=True)
    ctx = GEMMReduceScatterTensorParallelContext(rs_ctx=rs_ctx, output_dtype=output_dtype, rs_stream=rs_stream,
                                                 BLOCK_M=BLOCK_M, BLOCK_N=BLOCK_N if BLOCK_N != 256 else 0, BLOCK_K=BLOCK_K,
                                                 GROUP_M=GROUP_M, stages=stages)
    _symm.get_heap().barrier_all()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    return ctx


def torch_gemm_rs(a: torch.Tensor, b: torch.Tensor, group=None) -> torch.Tensor:
    """matmul + NCCL/gloo reduce-scatter: oracle and baseline (reference test_gemm_rs.py:38-51)."""
    return torch_reduce_scatter(torch.matmul(a, b.t()), group)


@launch_metadata("gemm_rs", lambda a, b, ctx, *_, **__: gemm_metadata(a.shape[0], b.shape[0], a.shape[1], a.element_size(), ctx.world_size,
                                                                     a.shape[0] * b.shape[0] * a.element_size() * (ctx.world_size - 1) // max(ctx.world_size, 1)))
def gemm_rs_op(input: torch.Tensor, weight: torch.Tensor, ctx: GEMMReduceScatterTensorParallelContext, persistent: bool = True,
               fuse_scatter: bool = True, output: Optional[torch.Tensor] = None, autotune: bool = False) -> torch.Tensor:
    _check_operands(input, weight)
    M, local_K = input.shape
    N = weight.shape[0]
    rs = ctx.rs_ctx
    W = rs.world_size
    assert N == rs.N and M % W == 0 and M <= rs.max_M
    if not rs.native:
        return torch_gemm_rs(input, weight)
    if output is None:
        output = torch.empty((M // W, N), dtype=ctx.output_dtype, device=input.device)
    if autotune:  # see allgather_gemm.ag_gemm: knobs from the contextual autotuner (protocol, tile kind)
        return _gemm_rs_tuned(input, weight, ctx, output)
    return _gemm_rs_launch(input, weight, ctx, output, ctx.variant, ctx.cta_group)


def _gemm_rs_launch(input, weight, ctx, output, variant, cta_group):
    M, local_K = input.shape
    N = weight.shape[0]
    rs = ctx.rs_ctx
    heap = _symm.get_heap()
    args = _gemm_args(input, weight, output, M, N, local_K, block_n=ctx.BLOCK_N, group_m=ctx.GROUP_M, cta_group=cta_group)
    st = rs.state
    f = _lib.FusedArgs(ctx=heap.ctx, mode=2, ws=rs.scatter_buf.data_ptr(), out=output.data_ptr(), flags=st.flags_ptr,
                       flags2=None, counters=ctx.counters.data_ptr(), phase=st.phase_ptr, done=st.done_ptr, num_comm_ctas=0,
                       ar_one_shot=0, use_mc=0, variant=variant,
                       ws_bytes=rs.scatter_buf.numel() * rs.scatter_buf.element_size(), flags_words=st.flags.numel())
    _lib.check(_lib.lib().tdist_gemm_fused(ctypes.byref(args), ctypes.byref(f), _lib.stream_ptr()), "gemm_rs")
    return output


def gemm_rs(a: torch.Tensor, b: torch.Tensor, ctx: GEMMReduceScatterTensorParallelContext, persistent: bool = True,
            fuse_scatter: bool = True, autotune: bool = False) -> torch.Tensor:
    """c[M/W, N] = reduce_scatter(a[M, K/W] @ b[N, K/W]^T)."""
    return gemm_rs_op(a, b, ctx, persistent, fuse_scatter, autotune=autotune)


from ...autotuner import tunable  # noqa: E402


@tunable("gemm_rs", {"variant": [0, 1, 2], "cta_group": [2, 1]},
         key=lambda input, weight, ctx, output=None, **_: ("gemm_rs", input.shape[0], weight.shape[0], input.shape[1],
                                                           ctx.rs_ctx.world_size, str(input.dtype)))
def _gemm_rs_tuned(input, weight, ctx, output, variant=0, cta_group=0):
    return _gemm_rs_launch(input, weight, ctx, output, variant, cta_group)
