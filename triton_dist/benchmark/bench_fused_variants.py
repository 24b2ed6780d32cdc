"""Sweep of the protocol / tile / push-CTA choices of AG+GEMM and GEMM+RS at one shape, in ONE launch: every
configuration is first checked against the fp32 NCCL oracle, then timed (device time, max over ranks, L2 flushed).
    torchrun --nproc-per-node 8 -m triton_dist.benchmark.bench_fused_variants [M N K] [--json out.json]
Default shape = BASELINE config 2 (Llama-3-8B MLP, M=8192 K=4096 N=14336). This is the measurement behind the defaults in
csrc/kernels/gemm.cu (reference counterpart: the autotune config lists of allgather_gemm.py:396-414)."""
import argparse
import json
import os

import torch

from triton_dist.kernels.nvidia.allgather_gemm import ag_gemm, create_ag_gemm_context, torch_ag_gemm
from triton_dist.kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs, torch_gemm_rs
from triton_dist.test.common import all_ok, check_close, finish, rand, setup
from triton_dist.utils import dist_print, perf_func_max_over_ranks

VAR = {1: "p2p", 2: "nvls"}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("M", type=int, nargs="?", default=8192)
    ap.add_argument("N", type=int, nargs="?", default=14336)
    ap.add_argument("K", type=int, nargs="?", default=4096)
    ap.add_argument("--iters", type=int, default=15)
    ap.add_argument("--json", default="")
    ap.add_argument("--quick", action="store_true", help="defaults only")
    ap.add_argument("--ablate", action="store_true", help="GEMM+RS timing ablations (TDIST_FUSED_DEBUG: results are wrong by design)")
    args = ap.parse_args()
    group, rank, world = setup()
    dt = torch.bfloat16
    M, N, K = args.M, args.N, args.K
    m_loc, n_loc = M // world, N // world
    rows = []
    all_good = True

    # ---------------------------------------------------------------- AG + GEMM
    A = rand((m_loc, K), dt)
    B = rand((n_loc, K), dt, 0.1)
    _, t_ref = perf_func_max_over_ranks(lambda: torch_ag_gemm(A, B), args.iters, 3, flush_l2=True)
    ref = torch_ag_gemm(A.float(), B.float())
    rows.append(dict(op="ag_gemm", cfg="torch NCCL+cuBLAS", ms=t_ref, ok=True))
    dist_print(f"ag_gemm  torch NCCL+cuBLAS: {t_ref:.4f} ms")
    # (variant, cta_group, push CTAs, row slices per block, K slabs)
    if args.quick or world == 1:
        ag_cfgs = [(0, 0, -1, 8, 1)]
    elif args.ablate:
        ag_cfgs = [(2, 2, 8, 8, 1), (2, 1, 8, 8, 1), (2, 2, 8, 8, 2), (2, 2, 4, 4, 1)]
    else:
        ag_cfgs = [(0, 0, -1, 8, 1), (2, 2, 8, 8, 1), (2, 2, 16, 16, 1), (2, 2, 16, 8, 1), (2, 1, 8, 8, 1), (2, 2, 8, 8, 2),
                   (1, 2, -1, 8, 1), (1, 1, -1, 8, 1)]
    for variant, cta_group, comm, slices, slabs in ag_cfgs:
        os.environ["TDIST_AG_SLICES"] = str(slices)
        os.environ["TDIST_AG_KSLABS"] = str(slabs)
        ctx = create_ag_gemm_context(A, B, rank, world, max_M=M, num_local_ranks=world, num_comm_sms=comm)
        ctx.variant, ctx.cta_group = variant, cta_group
        C = ag_gemm(A, B, ctx)
        good = all_ok(check_close(C, ref, dt, "ag_gemm"))
        _, t = perf_func_max_over_ranks(lambda: ag_gemm(A, B, ctx), args.iters, 3, flush_l2=True)
        name = f"{VAR.get(variant, 'auto')} cta_group={cta_group} push_ctas={comm} slices={slices} kslabs={slabs}"
        rows.append(dict(op="ag_gemm", cfg=name, ms=t, ok=good))
        dist_print(f"ag_gemm  {name:60s}: {t:.4f} ms  {t_ref / t:.2f}x  {'ok' if good else 'WRONG'}")
        all_good = all_good and good
        ctx.finalize()

    # ---------------------------------------------------------------- GEMM + RS  (down projection: K and N swap roles)
    k_loc = N // world
    A2 = rand((M, k_loc), dt)
    B2 = rand((K, k_loc), dt, 0.1)
    _, t_ref = perf_func_max_over_ranks(lambda: torch_gemm_rs(A2, B2), args.iters, 3, flush_l2=True)
    ref2 = torch_gemm_rs(A2.float(), B2.float())
    rows.append(dict(op="gemm_rs", cfg="torch cuBLAS+NCCL", ms=t_ref, ok=True))
    dist_print(f"gemm_rs  torch cuBLAS+NCCL: {t_ref:.4f} ms")
    rs_cfgs = [(0, 0, 0, 16)] if args.quick or world == 1 else [(0, 0, 0, 16), (2, 2, 0, 8), (2, 2, 0, 16), (2, 1, 0, 16), (1, 2, 0, 16), (1, 1, 0, 16)]
    if args.ablate:
        rs_cfgs = [(2, 2, 0, 16, d) for d in (0, 4, 8)] + [(2, 2, 0, 8, 0), (2, 1, 0, 8, 0)]
    for cfg in rs_cfgs:
        variant, cta_group, bn, unit_rows = cfg[:4]
        dbg = cfg[4] if len(cfg) > 4 else 0
        os.environ["TDIST_FUSED_DEBUG"] = str(dbg)
        os.environ["TDIST_RS_UNIT_ROWS"] = str(unit_rows)
        ctx = create_gemm_rs_context(M, K, rank, world, world, dt, BLOCK_N=bn)
        ctx.variant, ctx.cta_group = variant, cta_group
        C2 = gemm_rs(A2, B2, ctx)
        good = all_ok(dbg not in (0, 8) or check_close(C2, ref2, dt, "gemm_rs"))
        _, t = perf_func_max_over_ranks(lambda: gemm_rs(A2, B2, ctx), args.iters, 3, flush_l2=True)
        name = f"{VAR.get(variant, 'auto')} cta_group={cta_group} block_n={bn} unit_rows={unit_rows} debug={dbg}"
        rows.append(dict(op="gemm_rs", cfg=name, ms=t, ok=good))
        dist_print(f"gemm_rs  {name:60s}: {t:.4f} ms  {t_ref / t:.2f}x  {'ok' if good else 'WRONG'}")
        all_good = all_good and good
        ctx.finalize()
    for k in ("TDIST_AG_SLICES", "TDIST_AG_KSLABS", "TDIST_RS_UNIT_ROWS", "TDIST_FUSED_DEBUG"):
        os.environ.pop(k, None)
    if args.json and rank == 0:
        with open(args.json, "w") as f:
            json.dump(dict(world=world, M=M, N=N, K=K, rows=rows), f, indent=1)
    finish(all_good)
