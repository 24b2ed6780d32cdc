#!/usr/bin/env python
"""Headline benchmark (driver contract, see README/DESIGN).

Metric: aggregate TFLOP/s of one tensor-parallel MLP projection pair on the
Llama-3-8B shape named in BASELINE.json config 2 —
  step = AG+GEMM  (M=8192, K=4096,  N=14336 sharded over ranks)
       + GEMM+RS  (M=8192, K=14336 sharded over ranks, N=4096)
run at TP = number of GPUs, bf16, synthetic inputs, random weights. Total work is
fixed as GPUs are added (strong scaling). Device-timed with CUDA events around
every step, L2 flushed between steps, max over ranks.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nproc-per-node 8 ... bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl reference ...      # the unmodified reference, if it could be installed
    python bench.py --impl torch ...          # NCCL + cuBLAS through torch (the reference's own oracle path)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

M, K_HID, N_FFN = 8192, 4096, 14336
METRIC = "ag_gemm+gemm_rs aggregate TFLOP/s (Llama-3-8B MLP shape, TP=N)"


def reference_arm(args) -> None:
    """The reference needs its Triton fork (LLVM + NVIDIA toolchain downloads) and nvshmem4py; the offline
    install fails (DESIGN.md 'Reference arm'), so there is nothing to run."""
    ref = ROOT / "baseline" / "_ref"
    why = "reference not installable offline: python/setup.py downloads LLVM/toolchain (URLError), nvshmem4py absent"
    if ref.exists() and any(ref.iterdir()):
        why = "baseline/_ref present but triton_dist needs its patched Triton + NVSHMEM, which are not in this image"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))


class ClockSampler:
    def __init__(self, enabled: bool):
        self.proc = None
        self.path = Path(os.environ.get("TMPDIR", "/tmp")) / f"tdist_clocks_{os.getpid()}.csv"
        if enabled:
            try:
                q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
                     "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                     "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
                self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                             stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
            except Exception:
                self.proc = None

    def stop(self, n_gpus: int) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for ln in self.path.read_text().splitlines():
                f = [x.strip() for x in ln.split(",")]
                if len(f) < 9 or not f[0].isdigit() or int(f[0]) >= n_gpus:
                    continue
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                for name, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            self.path.unlink(missing_ok=True)
        except Exception:
            pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="tdist", choices=["tdist", "torch", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
        return
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist

    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", str(1))
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29555")
    from triton_dist.kernels.nvidia.allgather_gemm import ag_gemm, create_ag_gemm_context, torch_ag_gemm
    from triton_dist.kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs, torch_gemm_rs
    from triton_dist.utils import NVLINK_GBS_MEASURED, finalize_distributed, initialize_distributed, measured_peaks

    initialize_distributed(seed=1234)
    rank, world = dist.get_rank(), dist.get_world_size()
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torchrun for N>1)"
    dev = torch.device("cuda", torch.cuda.current_device())
    dtype = torch.bfloat16
    m_loc, n_loc, k_loc = M // world, N_FFN // world, N_FFN // world
    g = torch.Generator(device="cuda").manual_seed(1234 + rank)
    w_up = ((torch.rand(n_loc, K_HID, device=dev, generator=g) * 2 - 1) * 0.02).to(dtype)  # column-parallel [N/W, K]
    w_down = ((torch.rand(K_HID, k_loc, device=dev, generator=g) * 2 - 1) * 0.02).to(dtype)  # row-parallel [N, K/W]
    x_host = ((torch.rand(m_loc, K_HID, generator=torch.Generator().manual_seed(rank)) * 2 - 1)).to(dtype).pin_memory()
    x_dev = x_host.to(dev, non_blocking=True)
    ag_ctx = create_ag_gemm_context(x_dev, w_up, rank, world, max_M=M, num_local_ranks=world)
    rs_ctx = create_gemm_rs_context(M, K_HID, rank, world, world, dtype)
    h = torch.empty((M, n_loc), dtype=dtype, device=dev)
    y = torch.empty((m_loc, K_HID), dtype=dtype, device=dev)
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)
    res_host = torch.empty(8, dtype=torch.float32).pin_memory()

    if args.impl == "tdist":
        def step(x):
            ag_gemm(x, w_up, ag_ctx, out=h)
            from triton_dist.kernels.nvidia.gemm_reduce_scatter import gemm_rs_op
            return gemm_rs_op(h, w_down, rs_ctx, output=y)
        launches_per_step = 2
    else:
        def step(x):
            return torch_gemm_rs(torch_ag_gemm(x, w_up), w_down)
        launches_per_step = 0

    def timed(n_steps: int, e2e: bool):
        """Sum of per-step device times (ms); L2 flushed before every step, outside the timed span."""
        evs = []
        for _ in range(n_steps):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            if e2e:
                xin = x_host.to(dev, non_blocking=True)  # H2D of this step's input from pinned memory
                out = step(xin)
                res_host.copy_(out[0, :8].float(), non_blocking=True)  # D2H read of the step's result
            else:
                out = step(x_dev)
            e.record()
            evs.append((s, e))
        torch.cuda.synchronize()
        return sum(s.elapsed_time(e) for s, e in evs)

    def run(n_steps: int, e2e: bool) -> float:
        dist.barrier()
        torch.cuda.synchronize()
        ms = timed(n_steps, e2e)
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    run(args.warmup, False)
    sampler = ClockSampler(rank == 0)
    total_ms = run(args.steps, False)
    clocks = sampler.stop(world)
    run(max(args.warmup // 2, 2), True)
    e2e_ms = run(args.steps, True)

    # per-op breakdown (same timing rules), untimed w.r.t. the headline
    def op_ms(fn):
        for _ in range(3):
            fn()
        evs = []
        for _ in range(max(args.steps // 2, 5)):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            evs.append((s, e))
        torch.cuda.synchronize()
        ms = sum(s.elapsed_time(e) for s, e in evs) / len(evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    dist.barrier()
    ag_ms = op_ms((lambda: ag_gemm(x_dev, w_up, ag_ctx, out=h)) if args.impl == "tdist" else (lambda: torch_ag_gemm(x_dev, w_up)))
    rs_ms = op_ms((lambda: gemm_rs(h, w_down, rs_ctx)) if args.impl == "tdist" else (lambda: torch_gemm_rs(h, w_down)))
    base_ms = None
    if args.impl == "tdist":
        base_ms = op_ms(lambda: torch_gemm_rs(torch_ag_gemm(x_dev, w_up), w_down))

    flops_step = 2.0 * M * N_FFN * K_HID * 2  # whole job, both GEMMs
    ms_per_step = total_ms / args.steps
    value = flops_step / (ms_per_step * 1e-3) / 1e12
    pk = measured_peaks()
    per_gpu_flops = flops_step / 2 / world
    link_bytes = (world - 1) * m_loc * K_HID * 2
    roof_ms = max(per_gpu_flops / (pk["bf16_tflops"] * 1e9), link_bytes / (NVLINK_GBS_MEASURED * 1e6))
    out = {
        "metric": METRIC, "value": round(value, 2), "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 5), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic (uniform random activations, random-init weights)",
        "impl": args.impl,
        "config": {"model": "Llama-3-8B MLP projections (BASELINE.json config 2)", "global_batch": M, "seq_len": 1,
                   "M": M, "K": K_HID, "N": N_FFN, "parallelism": f"tp{world}", "l2": "flushed (256 MiB write) before every timed step",
                   "timing": "CUDA events per step, max over ranks"},
        "clocks": clocks,
        "e2e": {"value": round(flops_step / (e2e_ms / args.steps * 1e-3) / 1e12, 2), "unit": "TFLOP/s",
                "ms_per_step": round(e2e_ms / args.steps, 5), "h2d_bytes_per_step": x_host.numel() * 2, "d2h_bytes_per_step": 32},
        "gpu_launches": launches_per_step * args.steps,
        "breakdown": {
            "ag_gemm_ms": round(ag_ms, 5), "ag_gemm_tflops_per_gpu": round(per_gpu_flops / ag_ms / 1e9, 1),
            "ag_gemm_roofline_frac_of_" + pk["source"]: round(roof_ms / ag_ms, 3),
            "gemm_rs_ms": round(rs_ms, 5), "gemm_rs_tflops_per_gpu": round(per_gpu_flops / rs_ms / 1e9, 1),
            "gemm_rs_roofline_frac_of_" + pk["source"]: round(roof_ms / rs_ms, 3),
            "roofline_ms_per_op": round(roof_ms, 5),
            "torch_nccl_cublas_ms_per_step": None if base_ms is None else round(base_ms, 5),
            "speedup_vs_torch_nccl_cublas": None if base_ms is None else round(base_ms / ms_per_step, 3),
        },
    }
    if rank == 0:
        print(json.dumps(out))
    ag_ctx.finalize()
    rs_ctx.finalize()
    finalize_distributed()


if __name__ == "__main__":
    main()
