"""Tutorial 8 — overlapping GEMM with ReduceScatter, and what the measurements taught us about fences.
(Counterpart of the reference's tutorials/08-overlapping-gemm-reduce-scatter.py: a Triton producer GEMM that notifies
per-segment flags, a copy-engine scatter and a ring-reduce kernel on a second stream. Here: one CUDA kernel;
tutorial size in csrc/tutorials/overlap_tutorial.cu, production in csrc/kernels/gemm_sm100.cuh.)

    bash scripts/launch.sh tutorials/08-overlapping-gemm-reduce-scatter.py

The problem: a row-parallel linear layer. Rank r holds a K-slice A_r [M, K/W], B_r [N, K/W]; the layer output is the SUM of
the W partial products, and rank r only needs rows [r M/W, (r+1) M/W) of it (the token shard it owns).

STEP 0 — baseline: cuBLAS, then NCCL reduce-scatter of the full [M, N] partial. The links idle during the GEMM, the SMs
         idle during the collective.

STEP 1 — the obvious fusion ("scatter"): the GEMM epilogue stores each partial tile straight into the OWNER's memory (slot r
         of a [W][M/W][N] buffer there), a flag per block tells the owner, and the owner adds its W slots. This is what
         the reference does (FUSE_SCATTER) and what this library did in round 1. On an NVSwitch box it moves
         (W-1)/W of every partial over NVLink as stores out of the epilogue, and the add is a tail after the last tile.

STEP 2 — turn it around ("NVLS pull", ``tut_gemm_rs_kernel``): every rank keeps its partial tile in its OWN symmetric
         buffer and sends the owner only a flag. The owner then issues ``multimem.ld_reduce`` on the multicast alias of
         the tile: the switch fetches the W copies and adds them in fp32 on the way. Per rank 1/W of the data comes in,
         nothing is added on the SMs, and all ranks walk the tiles in the same owner-interleaved order (tile t belongs
         to owner t % W), so every owner has work all the time instead of only at the end.

STEP 3 — where do the reducers run? They are pure latency (a handful of loads in flight per thread saturate the link), so
         in the production kernel they are WARPS, not CTAs: warps 2 and 3 of every GEMM CTA were idle anyway (the CTA's
         TMA producer, MMA issuer and 4 epilogue warps do the GEMM), so no SM is taken from the tensor cores.

STEP 4 — what measuring taught us (8 x B200, M=8192 N=4096 K=14336/8; profiles/README.md has the table):
           * 0.230 ms  scatter protocol (step 1)                       NCCL + cuBLAS: 0.231 ms
           * 0.169 ms  NVLS pull, the epilogue thread publishes the flag with fence.acq_rel.sys + st.release.sys
           * 0.130 ms  same protocol, a third warp publishes with fence.acq_rel.GPU + a plain system-scope store
         A system-scope fence waits for everything the SM has in flight — including the reducer warp's NVLS loads, which
         take microseconds — and it stalled the epilogue, then the tensor core behind it. The partial tile lives in THIS
         GPU's memory and the switch reads it through this GPU's L2, so GPU scope is what the ordering needs.
         Ablations (timing only): without reducers the kernel takes 0.093 ms, i.e. the remaining 0.037 ms are flag
         latency and waiting for the slowest rank's tile.
"""
import ctypes
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist

from triton_dist import _lib
from triton_dist.kernels.common_ops import CollectiveState
from triton_dist.kernels.nvidia import create_gemm_rs_context, gemm_rs
from triton_dist.kernels.nvidia.gemm_reduce_scatter import torch_gemm_rs
from triton_dist.runtime import get_heap
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks, symm_tensor

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
if not torch.cuda.is_available():
    dist_print("this tutorial needs a GPU (the CPU plumbing backend has no device kernels)")
    finalize_distributed()
    raise SystemExit(0)
heap = get_heap()
M, N, Kl = 512 * world, 1024, 768  # global rows, output columns, this rank's slice of the reduction dim
torch.manual_seed(rank)
a = (torch.rand(M, Kl, device="cuda") * 2 - 1).bfloat16()
b = ((torch.rand(N, Kl, device="cuda") * 2 - 1) * 0.1).bfloat16()
ref = torch_gemm_rs(a.float(), b.float())  # step 0: fp32 oracle = matmul + NCCL reduce-scatter

# ---- step 2: the tutorial kernel
part = symm_tensor((2, M, N), torch.bfloat16, zero=False)  # every rank's partial product, two halves for consecutive calls
state = CollectiveState(world * ((M // world + 63) // 64) * ((N + 63) // 64) + 16)
heap.barrier_all()
out = torch.empty(M // world, N, device="cuda", dtype=torch.bfloat16)
lib = _lib.lib()
lib.tdist_tutorial_gemm_rs.restype = ctypes.c_int
lib.tdist_tutorial_gemm_rs.argtypes = ([ctypes.c_void_p] * 5 + [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 2 +
                                       [ctypes.c_int] * 4 + [ctypes.c_void_p])
flag_blocks = (M // world + 63) // 64   # constants of the buffers (see tutorial 7, step 2): the flag layout and the half-buffer stride


def tutorial_gemm_rs():
    _lib.check(lib.tdist_tutorial_gemm_rs(ctypes.byref(heap.ctx), a.data_ptr(), b.data_ptr(), out.data_ptr(), part.data_ptr(),
                                          part.numel() * part.element_size(), state.flags_ptr, flag_blocks, state.phase_ptr,
                                          state.done_ptr, M, N, Kl, 132, _lib.stream_ptr()), "tutorial_gemm_rs")
    return out


for call in range(3):
    tutorial_gemm_rs()
torch.cuda.synchronize()
err = (out.float() - ref).abs().max().item()
dist_print(f"rank {rank}: tutorial kernel (in-switch reduction: {bool(heap.multicast)}) max |err| = {err:.3e}", allowed_ranks="all")
assert err < 0.06 * ref.abs().max().item()

# ---- steps 3-4: the production op (reducer + publisher warps inside the tcgen05 GEMM CTAs)
ctx = create_gemm_rs_context(M, N, rank, world, world, torch.bfloat16)
y = gemm_rs(a, b, ctx)
assert (y.float() - ref).abs().max().item() < 0.06 * ref.abs().max().item()
_, t_tut = perf_func_max_over_ranks(tutorial_gemm_rs, 10, 3)
_, t_op = perf_func_max_over_ranks(lambda: gemm_rs(a, b, ctx), 10, 3)
_, t_ref = perf_func_max_over_ranks(lambda: torch_gemm_rs(a, b), 10, 3)
dist_print(f"M={M} N={N} K/W={Kl}: tutorial kernel (CUDA cores) {t_tut * 1e3:.1f} us | gemm_rs (tcgen05) {t_op * 1e3:.1f} us | "
           f"cuBLAS + NCCL {t_ref * 1e3:.1f} us")
ctx.finalize()
torch.cuda.synchronize()
heap.free(part)
state.free()
finalize_distributed()
