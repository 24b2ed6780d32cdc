"""Tutorial 5 — reduce-scatter inside one NVSwitch node: scatter-then-add, and reduction inside the switch.
(Counterpart of the reference's tutorials/05-intra-node-reduce-scatter.py, which scatters with copy engines into
per-rank slots and then runs a local reduction kernel, with ring variants for PCIe boxes. Kernels here:
``reduce_scatter_kernel`` and ``reduce_scatter_mc_kernel`` in csrc/kernels/collectives.cu.)

    bash scripts/launch.sh tutorials/05-intra-node-reduce-scatter.py

The problem: every rank holds a full partial result ``x_r [M, N]`` (e.g. the output of a row-parallel GEMM); rank r
needs ``sum_p x_p[r * M/W : (r+1) * M/W]`` — the sum over ranks of ITS row segment.

STEP 0 — baseline: ``torch.distributed.reduce_scatter_tensor`` (NCCL; on NVSwitch it uses NVLS itself).

STEP 1 — scatter, then add (``reduce_scatter_kernel``; this is the reference's structure with the copy engine replaced
by stores from the same kernel). The symmetric workspace holds W slots of one segment each:
        for every peer s:   store my segment s into slot[rank] of rank s          (W-1 segments leave over NVLink)
        per-CTA cross-rank barrier                                                (CTA b of every rank owns slice b)
        out = my own segment + slot[p] for all p != rank, summed in fp32          (reads W-1 segments from local HBM)
  Cross-rank synchronisation is PER CTA: CTA b of every rank handles the same slice of every segment, so "slice b has
  arrived from everybody" is a W-wide flag row per CTA; no grid-wide barrier, and early CTAs reduce while late CTAs are
  still storing. The workspace has two halves used by alternating calls, a FIXED distance apart (half the workspace,
  independent of this call's size), so back-to-back calls of different sizes never overlap and no trailing barrier is
  needed.

STEP 2 — let the switch add (``reduce_scatter_mc_kernel``). NVSwitch can execute a load on a multicast address as "fetch
this address from every GPU and return the SUM": ``multimem.ld_reduce.acc::f32.v4.bf16x2``. So:
        copy all my segments into the symmetric workspace (local HBM copy, same layout on every rank)
        per-CTA cross-rank barrier
        out[i] = multimem.ld_reduce(mc_ptr(workspace) + rank * segment + i)       (the switch reads W copies, adds in fp32)
  Per GPU only 1/W of the data comes IN (one already-reduced segment) instead of (W-1)/W, and no SM adds anything; each
  thread keeps 8 independent loads in flight (the link probe in profiles/README.md: 148 CTAs x 8 loads reach 714 GB/s).
  The library picks this kernel from 512 KB per segment when the heap has a multicast mapping; ``TDIST_RS_NO_NVLS=1``
  forces step 1.

STEP 3 — what the numbers say (8 x B200, profiles/README.md R2-4): 64 MB per rank: scatter-then-add 214 us, in-switch
156 us, NCCL 138 us. The stand-alone op still LOSES to NCCL by 10-30 % from 1 MB up: the local copy-in of the whole input
precedes the barrier (registering the producer's buffer in the heap, or chunking copy / barrier / reduce, removes it —
docs/roadmap.md). The real answer is not to run a stand-alone reduce-scatter at all: tutorials 6 and 8 fuse it into the
GEMM that produces ``x``, where the partial tiles are written into the symmetric buffer by the epilogue (no copy-in) and
reduced by idle warps while the tensor core works on later tiles — 0.129 ms against 0.231 ms for cuBLAS + NCCL.

The ring variants of the reference (``ring_reduce``, 2-D rings over PCIe / NUMA) exist as names: ``ring_reduce`` sums
the slots of a local buffer (``slot_reduce_kernel``), which is all that is left of a ring on a flat switch.
"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.reduce_scatter import (create_reduce_scater_2d_ctx, reduce_scatter_2d_op, ring_reduce,
                                                       torch_reduce_scatter)
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
gpu = torch.cuda.is_available()
dev = "cuda" if gpu else "cpu"
dtype = torch.bfloat16 if gpu else torch.float32
M, N = 1024 * world, 4096                                  # 8 MB per segment in bf16
torch.manual_seed(rank)
x = torch.randn(M, N, device=dev).to(dtype)

# ---- step 0: fp32 oracle
ref = torch_reduce_scatter(x.float())
scale = ref.abs().max().item()

# ---- steps 1 and 2 through the same op; the environment switch selects the kernel on a GPU with a multicast mapping
ctx = create_reduce_scater_2d_ctx(M, N, rank, world, world, dtype)
timings = {}
for name, no_nvls in (("scatter, then add", "1"), ("reduction in the switch", None)):
    if no_nvls is None:
        os.environ.pop("TDIST_RS_NO_NVLS", None)
    else:
        os.environ["TDIST_RS_NO_NVLS"] = no_nvls
    out = reduce_scatter_2d_op(x, ctx)
    err = (out.float() - ref).abs().max().item()
    dist_print(f"rank {rank}: {name:24s} [{M}, {N}] -> {tuple(out.shape)}  max |err| {err:.4f} (|ref| max {scale:.2f})",
               allowed_ranks="all")
    assert err <= 0.03 * scale + 1e-5                      # bf16 output of an fp32 sum over W terms
    if gpu:
        _, timings[name] = perf_func_max_over_ranks(lambda: reduce_scatter_2d_op(x, ctx), 20, 5)
os.environ.pop("TDIST_RS_NO_NVLS", None)

# ---- back-to-back calls of DIFFERENT sizes on one context: the fixed half-workspace parity stride at work
small = x[:M // 2]
for it in range(4):
    inp = x if it % 2 == 0 else small
    out = reduce_scatter_2d_op(inp, ctx)
    r = torch_reduce_scatter(inp.float())
    assert (out.float() - r).abs().max().item() <= 0.03 * scale + 1e-5
dist_print("alternating sizes on one context, no barrier in between: ok")

# ---- what is left of the ring: summing W local slots
slots = torch.randn(world, 256, N, device=dev).to(dtype)
acc = torch.empty(256, N, device=dev, dtype=dtype)
ring_reduce(slots.view(world * 256, N), acc, rank, world)
assert (acc.float() - slots.float().sum(0)).abs().max().item() <= 0.03 * slots.float().sum(0).abs().max().item() + 1e-5

if gpu:
    _, t_nccl = perf_func_max_over_ranks(lambda: torch_reduce_scatter(x), 20, 5)
    mb = M * N * x.element_size() / 2**20
    dist_print(f"{mb:.0f} MiB per rank: " + " | ".join(f"{k} {v * 1e3:.1f} us" for k, v in timings.items()) +
               f" | NCCL {t_nccl * 1e3:.1f} us")
    torch.cuda.synchronize()
ctx.finalize()
finalize_distributed()
