"""Tutorial 6 — the two-level ("inter-node") reduce-scatter, and its one-node form: the producer GEMM fused in.
(Counterpart of the reference's tutorials/06-inter-node-reduce-scatter.py: reduce inside the node over NVLink, exchange
per-node partial sums over InfiniBand with NVSHMEM puts, ring-reduce what arrives.)

    bash scripts/launch.sh tutorials/06-inter-node-reduce-scatter.py

WHY THE REFERENCE HAS TWO LEVELS. With n nodes of L GPUs a flat reduce-scatter would push (n*L - 1)/(n*L) of every
rank's data through its NIC. The reference therefore (1) scatters inside the node so that local rank l collects, for
every remote node, the node-local SUM of the segments that belong to that node's rank l (an intra-node reduce-scatter
per destination node, NVLink only), (2) sends ONE partial sum per destination node through the NIC (1/L of the flat
traffic), (3) ring-reduces the n partials that arrive. The stages are pipelined on three streams with per-stage
signals, and the producer GEMM notifies per tile so stage 1 starts before the GEMM has finished.

WHAT IS LEFT ON 8 x B200. There is no slow level: every peer is one NVLink hop away and the switch itself can add
(tutorial 5). What survives from that design, and matters more here than there, is the OVERLAP with the producer: a
reduce-scatter that starts when the GEMM has finished leaves the links idle for the whole GEMM and the tensor cores idle
for the whole collective. This tutorial walks from the two-kernel composition to the fused op:

STEP 0 — cuBLAS, then NCCL (the baseline every paper compares with).
STEP 1 — this library's GEMM, then its stand-alone reduce-scatter (two launches; same structure, our kernels).
STEP 2 — ``gemm_rs``: ONE kernel. The GEMM epilogue writes each finished partial tile into this rank's symmetric
         buffer and a publisher warp flags it to the tile's owner; a reducer warp in the OWNER's CTAs pulls the tile
         through ``multimem.ld_reduce`` (the switch adds the W partials) as soon as all W flags are in. All ranks walk
         the tiles owner-interleaved, so every rank has finished tiles to reduce from the first wave on, and the
         reduction of tile t overlaps the MMAs of tile t + 1. Tutorial 8 builds that kernel step by step.
STEP 3 — consecutive calls with DIFFERENT M on one context (prefill, then a shorter chunk): the workspace halves are a
         fixed distance apart and flags are epochs, so no barrier is needed between calls.

Measured on 8 x B200, M=8192 N=4096 K=14336 (profiles/README.md R2-1): step 0: 0.231 ms, step 2: 0.129 ms; the GEMM
alone takes 0.081 ms and NCCL's reduce-scatter of the 64 MB partial alone 0.138 ms: the fused kernel hides about two
thirds of the collective behind the tensor cores.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia import create_gemm_rs_context, gemm_persistent, gemm_rs
from triton_dist.kernels.nvidia.gemm_reduce_scatter import torch_gemm_rs
from triton_dist.kernels.nvidia.reduce_scatter import create_reduce_scater_2d_ctx, reduce_scatter_2d_op
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
gpu = torch.cuda.is_available()
dev = "cuda" if gpu else "cpu"
dtype = torch.bfloat16 if gpu else torch.float32

# ---- the two-level index arithmetic of the reference, for a pretend 2-node split of this job (nothing is sent)
L = max(world // 2, 1)
node, local = rank // L, rank % L
dist_print(f"rank {rank} = (node {node}, local {local}): stage 1 would reduce inside ranks {[node * L + l for l in range(L)]}, "
           f"stage 2 would send one partial to each of {[n * L + local for n in range(world // L) if n != node]}",
           allowed_ranks="all")

M, N, K = 1024 * world, 2048, 512 * world                 # row-parallel linear: every rank holds a K / W slice
torch.manual_seed(rank)
a = (torch.randn(M, K // world, device=dev) * 0.1).to(dtype)
b = (torch.randn(N, K // world, device=dev) * 0.1).to(dtype)
ref = torch_gemm_rs(a.float(), b.float())                 # step 0 in fp32 = the oracle
scale = ref.abs().max().item()
tol = 0.06 * scale if gpu else 1e-3 * scale + 1e-5

rs_ctx = create_gemm_rs_context(M, N, rank, world, world, dtype)
sc_ctx = create_reduce_scater_2d_ctx(M, N, rank, world, world, dtype)


def two_kernels(a_, b_):                                   # step 1
    return reduce_scatter_2d_op(gemm_persistent(a_, b_), sc_ctx)


def fused(a_, b_):                                         # step 2
    return gemm_rs(a_, b_, rs_ctx)


for name, fn in (("GEMM, then reduce-scatter (2 launches)", two_kernels), ("gemm_rs (1 launch)", fused)):
    err = (fn(a, b).float() - ref).abs().max().item()
    dist_print(f"{name:40s} max |err| {err:.4f} (|ref| max {scale:.2f})")
    assert err <= tol

# ---- step 3: alternate two problem sizes on the same contexts, no barrier and no host sync in between
a_small = a[:M // 2].contiguous()
ref_small = torch_gemm_rs(a_small.float(), b.float())
outs = []
for it in range(6):
    outs.append(fused(a, b) if it % 2 == 0 else fused(a_small, b))
for it, o in enumerate(outs):
    r = ref if it % 2 == 0 else ref_small
    assert (o.float() - r).abs().max().item() <= tol, f"call {it}"
dist_print("six back-to-back calls alternating M and M / 2 on one context: ok")

if gpu:
    _, t0 = perf_func_max_over_ranks(lambda: torch_gemm_rs(a, b), 20, 5)
    _, t1 = perf_func_max_over_ranks(lambda: two_kernels(a, b), 20, 5)
    _, t2 = perf_func_max_over_ranks(lambda: fused(a, b), 20, 5)
    _, tg = perf_func_max_over_ranks(lambda: gemm_persistent(a, b), 20, 5)
    dist_print(f"M={M} N={N} K/W={K // world}: cuBLAS + NCCL {t0 * 1e3:.1f} us | our GEMM + our reduce-scatter {t1 * 1e3:.1f} us | "
               f"fused {t2 * 1e3:.1f} us | (GEMM alone {tg * 1e3:.1f} us)")
    torch.cuda.synchronize()
rs_ctx.finalize()
sc_ctx.finalize()
finalize_distributed()
