#!/usr/bin/env python
"""Headline benchmark (driver contract, see README/DESIGN).

Metric: aggregate TFLOP/s of one tensor-parallel MLP projection pair on the Llama-3-8B shape named in BASELINE.json
config 2 —
  step = AG+GEMM  (M=8192, K=4096,  N=14336 sharded over ranks)
       + GEMM+RS  (M=8192, K=14336 sharded over ranks, N=4096)
run at TP = number of GPUs, bf16, synthetic inputs, random weights. Total work is fixed as GPUs are added (strong
scaling). Device-timed with CUDA events around every step, L2 flushed between steps, max over ranks.
The JSON line also carries the second half of the BASELINE.json metric, Qwen3-32B decode ms/token at TP=N (bs=1,
ctx=512, random-init weights, one megakernel launch per token), under "decode".

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nproc-per-node 8 ... bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl reference ...      # the unmodified reference, if it could be installed
    python bench.py --impl torch ...          # NCCL + cuBLAS through torch (the reference's own oracle path)
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

M, K_HID, N_FFN = 8192, 4096, 14336
METRIC = "ag_gemm+gemm_rs aggregate TFLOP/s (Llama-3-8B MLP shape, TP=N)"


def reference_arm(args) -> None:
    """The reference needs its Triton fork (LLVM + NVIDIA toolchain downloads) and nvshmem4py; the offline install
    fails, so there is nothing to run."""
    ref = ROOT / "baseline" / "_ref"
    why = "reference not installable offline: python/setup.py downloads LLVM/toolchain (URLError), nvshmem4py absent"
    if ref.exists() and any(ref.iterdir()):
        why = "baseline/_ref present but triton_dist needs its patched Triton + NVSHMEM, which are not in this image"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))


class ClockSampler:
    """NVML poller (SM clock + throttle reasons of the GPUs in use) running in a thread while the timed region executes;
    a 5 ms period gives samples even for regions of a few tens of milliseconds."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, n_gpus: int, enabled: bool):
        self.samples, self.max_clock, self.reasons = [], None, set()
        self._stop = threading.Event()
        self._thread = None
        self.nv = None
        if not enabled:
            return
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.handles = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(n_gpus)]
            self.max_clock = max(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM) for h in self.handles)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                for h in self.handles:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                    for bit, name in self.REASONS.items():
                        if mask & bit:
                            self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self) -> dict:
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_clock, "reasons": ["unsampled"], "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_clock, "reasons": sorted(self.reasons), "samples": len(s)}


def decode_section(rank: int, world: int, steps: int):
    """Qwen3-32B decode latency at TP=world: bs=1, ctx=512, megakernel (one launch per token)."""
    import torch
    import torch.distributed as dist

    from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
    from triton_dist.models import DenseLLM, ModelConfig
    from triton_dist.utils import measured_peaks, perf_func_max_over_ranks

    cfg = ModelConfig(model_name="Qwen/Qwen3-32B", max_length=1024, dtype=torch.bfloat16, rank=rank, world_size=world)
    dense = DenseLLM(cfg, None)
    builder = ModelBuilder(rank, world, world)
    mega = Qwen3Model(1, cfg, builder, dense=dense)
    mega.set_kv_len(512, fill_random=True)
    ids = torch.full((1,), 1225, dtype=torch.long, device="cuda")

    def step():
        mega.kv_cache.set_position(512)
        return mega.mega_forwrad(ids)

    _, ms = perf_func_max_over_ranks(step, steps, 3)
    wbytes = dense.lm_head.numel() * 2 + sum(w.numel() * 2 for l in dense.layers for w in (l.attn.wqkv, l.attn.wo, l.mlp.gate_up_proj, l.mlp.down_proj))
    pk = measured_peaks()
    floor = wbytes / (pk["hbm_gbs"] * 1e6)
    info = builder.info()
    out = {"model": "Qwen3-32B (random init)", "tp": world, "bsz": 1, "ctx": 512, "ms_per_token": round(ms, 4),
           "weight_bytes_per_token_per_rank": wbytes, "hbm_floor_ms_of_" + pk["source"]: round(floor, 4),
           "fraction_of_hbm_floor": round(floor / ms, 3), "launches_per_token": 1, "graph": info,
           "reference_published_ms_8xH800": 7.41 if world == 8 else None}
    try:  # per-op-type timeline of one token (in-kernel task stamps), outside the timed region
        builder.enable_profiling(True)
        step()
        torch.cuda.synchronize()
        out["task_profile"] = builder.profile_summary()
        builder.enable_profiling(False)
    except Exception as e:  # noqa: BLE001
        out["task_profile"] = f"unavailable: {e}"
    mega.finalize()
    del dense
    torch.cuda.empty_cache()
    return out


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="tdist", choices=["tdist", "torch", "reference"])
    ap.add_argument("--no-decode", action="store_true", help="skip the Qwen3-32B decode section")
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
    from triton_dist.kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs, gemm_rs_op, torch_gemm_rs
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
            return gemm_rs_op(h, w_down, rs_ctx, output=y)
        launches_per_step = 2
    else:
        def step(x):
            return torch_gemm_rs(torch_ag_gemm(x, w_up), w_down)
        launches_per_step = 0

    def timed_device(n_steps: int) -> float:
        """Sum of per-step device times (ms); L2 flushed before every step, outside the timed span."""
        evs = []
        for _ in range(n_steps):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            step(x_dev)
            e.record()
            evs.append((s, e))
        torch.cuda.synchronize()
        return sum(s.elapsed_time(e) for s, e in evs)

    # end-to-end: every step copies its input from pinned host memory and reads a result back. The copies run on a
    # copy stream into two alternating device buffers so that the H2D of step i+1 overlaps the compute of step i
    # (what a serving loop does); the timed span covers all copies, all compute and all read-backs of the K steps.
    copy_stream = torch.cuda.Stream()
    x_bufs = [torch.empty_like(x_dev), torch.empty_like(x_dev)]

    def timed_e2e(n_steps: int) -> float:
        cur = torch.cuda.current_stream()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        copy_stream.wait_stream(cur)
        start.record(cur)
        copy_stream.wait_event(start)
        for i in range(n_steps):
            b = i & 1
            with torch.cuda.stream(copy_stream):
                if i >= 2:
                    copy_stream.wait_event(consumed[b])  # buffer b is free once step i-2 has read it
                x_bufs[b].copy_(x_host, non_blocking=True)  # H2D of this step's input
                ready[b].record(copy_stream)
            cur.wait_event(ready[b])
            out = step(x_bufs[b])
            consumed[b].record(cur)
            res_host.copy_(out[0, :8].float(), non_blocking=True)  # D2H read of the step's result
        stop.record(cur)
        torch.cuda.synchronize()
        return start.elapsed_time(stop)

    def run(fn, n_steps: int) -> float:
        dist.barrier()
        torch.cuda.synchronize()
        ms = fn(n_steps)
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(world, rank == 0)
    run(timed_device, args.warmup)
    sampler.start()
    total_ms = run(timed_device, args.steps)
    clocks = sampler.stop()
    run(timed_e2e, max(args.warmup // 2, 3))
    e2e_ms = run(timed_e2e, args.steps)

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
                "ms_per_step": round(e2e_ms / args.steps, 5), "h2d_bytes_per_step": x_host.numel() * 2, "d2h_bytes_per_step": 32,
                "note": "H2D of step i+1 overlaps compute of step i (copy stream, two device buffers)"},
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
    ag_ctx.finalize()
    rs_ctx.finalize()
    if args.impl == "tdist" and not args.no_decode and os.environ.get("TDIST_BENCH_DECODE", "1") != "0":
        try:
            out["decode"] = decode_section(rank, world, max(args.steps, 10))
        except Exception as e:  # the headline number must survive a failure of the secondary section
            out["decode"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    if rank == 0:
        print(json.dumps(out))
    finalize_distributed()


if __name__ == "__main__":
    main()
