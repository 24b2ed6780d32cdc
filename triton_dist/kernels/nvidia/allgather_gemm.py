"""AllGather + GEMM (tensor-parallel column linear with sequence-sharded input).

API of the reference's ``kernels/nvidia/allgather_gemm.py`` (context :417-531,
``ag_gemm`` :534-653, ``gemm_persistent`` / ``gemm_non_persistent`` :656-720).

B200 implementation: ONE launch of the tcgen05 GEMM engine
(csrc/kernels/gemm_sm100.cuh, mode AG). A few CTAs push this rank's A shard
into every rank's gather workspace — with NVLS ``multimem.st`` (each byte leaves
the GPU once, the switch replicates it; all ranks push their 128-row blocks in
the same order) or, without a multicast mapping, with bulk async copies to each
peer — and release flags per block; the remaining CTAs run the persistent
TMA -> tcgen05 -> TMEM GEMM over the blocks in arrival order (own shard first,
read straight from the caller's tensor). No copy engine, no second stream, no
host-side barrier: the workspace is double buffered by call parity (the two
halves are a fixed distance apart, whatever the size of a call) and flags are
monotone, so back-to-back calls of any sizes and CUDA graphs are safe.
"""
from __future__ import annotations

import ctypes
import dataclasses
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ...utils import gemm_metadata, launch_metadata
from ..common_ops import CollectiveState
from .allgather import AllGatherMethod, get_auto_all_gather_method


def _gemm_args(a: torch.Tensor, b: torch.Tensor, c: torch.Tensor, M: int, N: int, K: int, block_n: int = 0, group_m: int = 0,
               num_sms: int = 0, transpose_out: bool = False, cta_group: int = 0) -> _lib.GemmArgs:
    return _lib.GemmArgs(a=a.data_ptr(), b=b.data_ptr(), c=c.data_ptr(), M=M, N=N, K=K, lda=a.stride(0), ldb=b.stride(0),
                         ldc=c.stride(0), dtype=_lib.dtype_code(a.dtype), block_n=block_n, group_m=group_m, num_sms=num_sms,
                         transpose_out=int(transpose_out), block_expert=None, num_experts=0, cta_group=cta_group)


def _check_operands(a: torch.Tensor, b: torch.Tensor) -> None:
    assert a.dim() == 2 and b.dim() == 2 and a.shape[1] == b.shape[1], \
        f"B must be [N, K] (nn.Linear layout) with K={a.shape[1]}, got {tuple(b.shape)}"
    assert a.dtype == b.dtype, f"dtype mismatch: A {a.dtype}, B {b.dtype}"
    assert a.stride(1) == 1 and b.stride(1) == 1, "operands must be K-major (row-major [rows, K])"


@launch_metadata("gemm", lambda a, b, *_, **__: gemm_metadata(a.shape[0], b.shape[0], a.shape[1], a.element_size()))
def gemm_persistent(a: torch.Tensor, b: torch.Tensor, ctx=None, out: Optional[torch.Tensor] = None, block_n: int = 0,
                    num_sms: int = 0, cta_group: int = 0, **_ignored) -> torch.Tensor:
    """C = A @ B^T on the tcgen05 engine (single GPU; reference :656-688)."""
    _check_operands(a, b)
    M, K = a.shape
    N = b.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=a.dtype, device=a.device)
    if not _lib.use_native():
        torch.matmul(a, b.t(), out=out)
        return out
    # skinny problems: swap operands so the weight fills the 128-row MMA dimension
    if M <= 64 and N >= 128 and M % 8 == 0:
        args = _gemm_args(b, a, out, N, M, K, block_n=block_n, num_sms=num_sms, transpose_out=True)
    else:
        args = _gemm_args(a, b, out, M, N, K, block_n=block_n, num_sms=num_sms, cta_group=cta_group)
    _lib.check(_lib.lib().tdist_gemm(ctypes.byref(args), _lib.stream_ptr()), "gemm")
    return out


def gemm_non_persistent(a: torch.Tensor, b: torch.Tensor, ctx=None, **kw) -> torch.Tensor:
    """The reference's one-tile-per-CTA schedule; on B200 the persistent engine covers it."""
    return gemm_persistent(a, b, ctx, **kw)


@dataclasses.dataclass
class AllGatherGEMMTensorParallelContext:
    max_M: int
    N_per_rank: int
    K: int
    tensor_dtype: torch.dtype
    rank: int
    num_ranks: int
    num_local_ranks: int
    BLOCK_M: int = 128
    BLOCK_N: int = 0  # 0 = wave-quantisation heuristic (csrc/kernels/gemm.cu:pick_block_n)
    BLOCK_K: int = 64
    GROUP_SIZE_M: int = 8
    stages: int = 4
    num_comm_sms: int = -1  # CTAs that push the local shard to the peers (-1 = chosen so that no GEMM wave is added)
    cta_group: int = 0  # 0 auto, 1 single-CTA tiles, 2 CTA pairs
    variant: int = 0  # 0 auto (NVLS broadcast push when the heap has a multicast mapping), 1 point-to-point, 2 NVLS
    all_gather_method: AllGatherMethod = AllGatherMethod.Auto
    for_correctness: bool = False
    ag_intranode_stream: Optional["torch.cuda.Stream"] = None
    ag_internode_stream: Optional["torch.cuda.Stream"] = None
    symm_workspace: Optional[torch.Tensor] = dataclasses.field(default=None, init=False)
    symm_workspaces: List[torch.Tensor] = dataclasses.field(default_factory=list, init=False)
    state: Optional[CollectiveState] = dataclasses.field(default=None, init=False)

    def __post_init__(self):
        assert self.num_ranks % self.num_local_ranks == 0
        self.is_multinode = self.num_ranks > self.num_local_ranks
        if self.is_multinode:
            raise NotImplementedError("this framework targets one NVSwitch node (num_ranks == num_local_ranks)")
        self.n_nodes, self.node_rank, self.local_rank = 1, 0, self.rank
        heap = _symm.get_heap()
        self.native = isinstance(heap, _symm.SymmetricHeap)
        if self.native:
            # [2 parity][max_M][K] gather workspace + per-(src, 128-row block) arrival flags
            self.symm_workspace = heap.alloc((2, self.max_M, self.K), self.tensor_dtype, zero=False)
            self.symm_workspaces = heap.peer_views(self.symm_workspace)
            # 128 flag words per (source rank, 128-row block): K slabs x push CTAs sharing the block
            self.flags_words = 128 * (self.num_ranks * ((self.max_M // self.num_ranks + 127) // 128) + 16)
            self.state = CollectiveState(self.flags_words)
            self.max_gemm_sm = torch.cuda.get_device_properties("cuda").multi_processor_count
            heap.barrier_all()
            torch.cuda.synchronize()

    def finalize(self):
        if self.native and self.symm_workspace is not None:
            torch.cuda.synchronize()
            heap = _symm.get_heap()
            heap.free(self.symm_workspace)
            self.state.free()
            self.symm_workspace = None

    finailize = finalize  # the reference's spelling (allgather_gemm.py:483), used by its layers


def create_ag_gemm_context(tensor_A: torch.Tensor, tensor_B: torch.Tensor, rank: int, num_ranks: int, max_M: int,
                           num_local_ranks: int = 8, BLOCK_M: int = 128, BLOCK_N: int = 0, BLOCK_K: int = 64, stages: int = 4,
                           ag_intranode_stream=None, ag_internode_stream=None, for_correctness: bool = False,
                           num_comm_sms: int = -1) -> AllGatherGEMMTensorParallelContext:
    """tensor_A: local [M_per_rank, K]; tensor_B: local [N_per_rank, K]; max_M >= M_per_rank * num_ranks."""
    _check_operands(tensor_A, tensor_B)
    num_local_ranks = min(num_local_ranks, num_ranks)
    return AllGatherGEMMTensorParallelContext(
        max_M=max_M, N_per_rank=tensor_B.shape[0], K=tensor_A.shape[1], tensor_dtype=tensor_A.dtype, rank=rank,
        num_ranks=num_ranks, num_local_ranks=num_local_ranks, BLOCK_M=BLOCK_M, BLOCK_N=BLOCK_N if BLOCK_N != 256 else 0,
        BLOCK_K=BLOCK_K, stages=stages, ag_intranode_stream=ag_intranode_stream, ag_internode_stream=ag_internode_stream,
        for_correctness=for_correctness, num_comm_sms=num_comm_sms,
        all_gather_method=get_auto_all_gather_method(num_ranks, num_local_ranks))


def torch_ag_gemm(a: torch.Tensor, b: torch.Tensor, group=None) -> torch.Tensor:
    """NCCL/gloo all-gather + matmul: the oracle and the baseline (reference test_ag_gemm.py:117-128)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return torch.matmul(a, b.t())
    full = torch.empty((a.shape[0] * world, a.shape[1]), dtype=a.dtype, device=a.device)
    dist.all_gather_into_tensor(full, a.contiguous(), group=group)
    return torch.matmul(full, b.t())


@launch_metadata("ag_gemm", lambda a, b, ctx, *_, **__: gemm_metadata(a.shape[0] * ctx.num_ranks, b.shape[0], a.shape[1], a.element_size(), ctx.num_ranks,
                                                                     (ctx.num_ranks - 1) * a.numel() * a.element_size()))
def ag_gemm(a: torch.Tensor, b: torch.Tensor, ctx: AllGatherGEMMTensorParallelContext, persistent: bool = True,
            autotune: bool = False, straggler_option=None, use_cooperative: bool = True,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """C[M, N_per_rank] = all_gather(a)[M, K] @ b[N_per_rank, K]^T, M = M_per_rank * num_ranks."""
    _check_operands(a, b)
    M_per_rank, K = a.shape
    N = b.shape[0]
    M = M_per_rank * ctx.num_ranks
    assert M <= ctx.max_M and K == ctx.K, f"context built for max_M={ctx.max_M}, K={ctx.K}; got M={M}, K={K}"
    if not ctx.native:
        return torch_ag_gemm(a, b)
    if out is None:
        out = torch.empty((M, N), dtype=a.dtype, device=a.device)
    if straggler_option is not None and straggler_option[0] == ctx.rank:
        torch.cuda._sleep(int(straggler_option[1]))
    if ctx.for_correctness:  # shake out ordering bugs: random skew between ranks
        torch.cuda._sleep(int(torch.randint(0, 2_000_000, (1,)).item()))
    if autotune:
        # launch knobs come from the contextual autotuner: inside contextual_autotune(...) every combination is tried (the
        # winner = smallest MAX-over-ranks time is cached per shape), outside it the cached winner (or the defaults) runs
        return _ag_gemm_tuned(a, b, ctx, out)
    return _ag_gemm_launch(a, b, ctx, out, ctx.variant, ctx.cta_group, ctx.num_comm_sms)


def _ag_gemm_launch(a, b, ctx, out, variant, cta_group, num_comm_sms):
    M_per_rank, K = a.shape
    N = b.shape[0]
    M = M_per_rank * ctx.num_ranks
    heap = _symm.get_heap()
    args = _gemm_args(a, b, out, M, N, K, block_n=ctx.BLOCK_N, group_m=ctx.GROUP_SIZE_M, cta_group=cta_group)
    st = ctx.state
    f = _lib.FusedArgs(ctx=heap.ctx, mode=1, ws=ctx.symm_workspace.data_ptr(), out=None, flags=st.flags_ptr, flags2=None,
                       counters=None, phase=st.phase_ptr, done=st.done_ptr, num_comm_ctas=num_comm_sms, ar_one_shot=0,
                       use_mc=0, variant=variant, ws_bytes=ctx.symm_workspace.numel() * ctx.symm_workspace.element_size(),
                       flags_words=ctx.flags_words)
    _lib.check(_lib.lib().tdist_gemm_fused(ctypes.byref(args), ctypes.byref(f), _lib.stream_ptr()), "ag_gemm")
    return out


def _ag_key(a, b, ctx, out=None, **_):
    return ("ag_gemm", a.shape[0] * ctx.num_ranks, b.shape[0], a.shape[1], ctx.num_ranks, str(a.dtype))


# the space the reference tunes its consumer GEMM over (allgather_gemm.py:396-414: tile shapes / stages / warps) becomes,
# for ahead-of-time compiled kernels: protocol (1 point-to-point, 2 NVLS; 0 = built-in choice), tile kind (CTA pair or
# single CTA) and number of push CTAs. Without a multicast mapping an NVLS entry runs the built-in choice instead.
from ...autotuner import tunable  # noqa: E402

_AG_SPACE = {"variant": [0, 1, 2], "cta_group": [2, 1], "num_comm_sms": [-1, 8, 16]}


@tunable("ag_gemm", _AG_SPACE, key=_ag_key)
def _ag_gemm_tuned(a, b, ctx, out, variant=0, cta_group=0, num_comm_sms=-1):
    if variant == 2 and not _symm.get_heap().multicast:
        variant = 0
    return _ag_gemm_launch(a, b, ctx, out, variant, cta_group, num_comm_sms)
