"""Tutorial 7 — overlapping AllGather with GEMM, built up in four steps.
(Counterpart of the reference's tutorials/07-overlapping-allgather-gemm.py, which builds a Triton consumer GEMM that
``dl.wait``s on per-rank flags while the copy engine gathers. Here the steps are CUDA: the tutorial kernel lives in
csrc/tutorials/overlap_tutorial.cu, the production kernel in csrc/kernels/gemm_sm100.cuh.)

    bash scripts/launch.sh tutorials/07-overlapping-allgather-gemm.py          # all GPUs of the node
    python tutorials/07-overlapping-allgather-gemm.py                          # one GPU: the protocol degenerates

The problem: a column-parallel linear layer whose input is sharded on tokens (Megatron sequence parallelism).
Rank r holds A_r [M/W, K] and its slice of the weight B_r [N/W, K]; it needs C_r = all_gather(A) @ B_r^T.

STEP 0 — the baseline everybody starts from: NCCL all-gather, then cuBLAS. The GEMM cannot start before the LAST byte
         of the gather has arrived, and the links idle while the GEMM runs.

STEP 1 — symmetric memory. Every rank allocates the same buffers at the same offsets of one heap that is mapped into
         all processes (triton_dist.utils.symm_tensor; csrc/runtime/symm_heap.cc: cuMemCreate + fd exchange + one VA
         range with rank p at base + p * bytes). A kernel can therefore store into a peer's copy of a buffer
         (``symm_at(ctx, ptr, peer)``) or into ALL copies at once through the NVLS multicast alias (``mc_ptr``).

STEP 2 — flags instead of barriers (csrc/device/symm.cuh). The producer of a block does
             payload stores; __syncthreads(); one thread: fence(); notify(flag, peer, epoch)
         and a consumer of that block
             one thread: wait_ge_sys(flag, epoch); __syncthreads(); loads
         ``epoch`` grows by 2 with every call of the op and is never reset, and the gather workspace has two halves that
         calls use alternately: a rank that is already in call n+1 writes the OTHER half, so no barrier is needed between
         calls, and the whole op is one CUDA-graph-capturable launch. Two rules make that safe when calls differ in
         size: the halves are HALF THE WORKSPACE apart and a flag's position depends only on (rank, block) and the size
         of the flag array, never on the rows of the current call. The CPU protocol simulator (csrc/sim, run under
         ThreadSanitizer by tests/test_cpu_protocol_sim.py) executes this very kernel source with 2-8 simulated ranks and
         shows the race, and the lost wake-up, that either shortcut would cause.

STEP 3 — one kernel, two kinds of CTAs (``tut_ag_gemm_kernel``). The first few CTAs push this rank's shard, one 64-row
         block after the other, to every rank and flag each block. All other CTAs compute output tiles in the order the
         data arrives: own shard first (no wait at all), then block 0 of every remote shard, block 1, ... Each tile waits
         only for the one block it reads. With a multicast mapping a block leaves the GPU ONCE (``multimem.st``) and the
         switch replicates it — at 8 ranks that is 8 MiB out instead of 56 MiB for the Llama-3-8B shape.

STEP 4 — the same structure at full speed (``ag_gemm``): the inner loops become TMA loads into a 6-stage shared-memory
         ring, ``tcgen05.mma`` with the accumulator in tensor memory, CTA pairs computing 256-row tiles, and the waiting
         is done by the TMA producer warp, lanes polling the flags of a block in parallel. Measured on 8 x B200 for
         M=8192 K=4096 N=14336: 0.149 ms against 0.231 ms for NCCL + cuBLAS (profiles/README.md).
"""
import ctypes
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist

from triton_dist import _lib
from triton_dist.kernels.common_ops import CollectiveState
from triton_dist.kernels.nvidia import ag_gemm, create_ag_gemm_context
from triton_dist.kernels.nvidia.allgather_gemm import torch_ag_gemm
from triton_dist.runtime import get_heap
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks, symm_tensor

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
if not torch.cuda.is_available():
    dist_print("this tutorial needs a GPU (the CPU plumbing backend has no device kernels)")
    finalize_distributed()
    raise SystemExit(0)
heap = get_heap()
rows, K, N = 512, 1024, 768  # per-rank rows of A, reduction dim, per-rank columns
torch.manual_seed(rank)
a = (torch.rand(rows, K, device="cuda") * 2 - 1).bfloat16()
b = ((torch.rand(N, K, device="cuda") * 2 - 1) * 0.1).bfloat16()

# ---- step 0: the oracle and baseline
ref = torch_ag_gemm(a.float(), b.float())

# ---- steps 1-3: symmetric workspace + flags + the tutorial kernel
ws = symm_tensor((2, world * rows, K), torch.bfloat16, zero=False)  # two halves, used alternately by consecutive calls
state = CollectiveState(world * ((rows + 63) // 64) + 16)            # flags [W][blocks] + the (phase, done) counters
heap.barrier_all()                                                   # everybody has zeroed its flags before anybody pushes
c = torch.empty(world * rows, N, device="cuda", dtype=torch.bfloat16)
lib = _lib.lib()
lib.tdist_tutorial_ag_gemm.restype = ctypes.c_int
lib.tdist_tutorial_ag_gemm.argtypes = ([ctypes.c_void_p] * 5 + [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 2 +
                                       [ctypes.c_int] * 5 + [ctypes.c_void_p])
flag_blocks = (rows + 63) // 64   # a constant of the flag array, like the half-workspace stride: never a size of one call


def tutorial_ag_gemm():
    _lib.check(lib.tdist_tutorial_ag_gemm(ctypes.byref(heap.ctx), a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(),
                                          ws.numel() * ws.element_size(), state.flags_ptr, flag_blocks, state.phase_ptr,
                                          state.done_ptr, rows, N, K, 4, 132, _lib.stream_ptr()), "tutorial_ag_gemm")
    return c


for call in range(3):  # several calls back to back: no barrier, no reset between them
    tutorial_ag_gemm()
torch.cuda.synchronize()
err = (c.float() - ref).abs().max().item()
dist_print(f"rank {rank}: tutorial kernel (multicast push: {bool(heap.multicast)}) max |err| = {err:.3e}", allowed_ranks="all")
assert err < 0.06 * ref.abs().max().item()

# ---- step 4: the production op on the same problem, and its time against the baseline
ctx = create_ag_gemm_context(a, b, rank, world, max_M=world * rows, num_local_ranks=world)
c2 = ag_gemm(a, b, ctx)
assert (c2.float() - ref).abs().max().item() < 0.06 * ref.abs().max().item()
_, t_tut = perf_func_max_over_ranks(tutorial_ag_gemm, 10, 3)
_, t_op = perf_func_max_over_ranks(lambda: ag_gemm(a, b, ctx), 10, 3)
_, t_ref = perf_func_max_over_ranks(lambda: torch_ag_gemm(a, b), 10, 3)
dist_print(f"M={world * rows} N/W={N} K={K}: tutorial kernel (CUDA cores) {t_tut * 1e3:.1f} us | ag_gemm (tcgen05) {t_op * 1e3:.1f} us | "
           f"NCCL + cuBLAS {t_ref * 1e3:.1f} us")
ctx.finalize()
torch.cuda.synchronize()
heap.free(ws)
state.free()
finalize_distributed()
