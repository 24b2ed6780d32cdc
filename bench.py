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

M, K_HID, N_FFN = 8192, 4096, 14336
METRIC = "ag_gemm+gemm_rs aggregate TFLOP/s (Llama-3-8B MLP shape, TP=N)"
NVLINK_GBS = 770.0  # measured peer-copy rate per direction per GPU (B200_PROFILING.md)
REFERENCE_CLASS = ("reference TP_MLP.torch_ag_gemm + TP_MLP.torch_gemm_rs (layers/nvidia/tp_mlp.py:231-264, NCCL + cuBLAS): "
                   "the unmodified reference source; its Triton kernels cannot be built offline (DESIGN.md section 6)")


def measured_peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"], "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "source": "fallback"}


def reference_tree():
    """Where the unmodified reference package lives: baseline/_ref (populated by scripts/install_reference.sh, travels
    to the GPU box) or the read-only checkout of the authoring container."""
    for cand in (ROOT / "baseline" / "_ref" / "triton_dist", Path("/root/reference/python/triton_dist")):
        if (cand / "layers" / "nvidia" / "tp_mlp.py").exists():
            return cand
    return None


def load_reference_tp_mlp(tree: Path):
    """Import the reference's layers/nvidia/tp_mlp.py UNMODIFIED. Its module-level imports of the Triton-compiled kernel
    modules (which need the reference's own Triton fork + NVSHMEM, not buildable offline) resolve to empty placeholder
    modules, so only its torch / NCCL methods can run: touching anything else raises. Nothing of this repository is
    importable in this process (the repo root is not on sys.path and `triton_dist` is claimed by the placeholders)."""
    import importlib.abc
    import importlib.machinery
    import importlib.util
    import types

    class _Placeholder:
        def __init__(self, name):
            self._name = name

        def __call__(self, *a, **k):
            raise RuntimeError(f"reference symbol {self._name} needs the reference's Triton/NVSHMEM build (unavailable offline)")

        def __getattr__(self, item):
            if item.startswith("__"):
                raise AttributeError(item)
            return _Placeholder(f"{self._name}.{item}")

    class _StubModule(types.ModuleType):
        __path__ = []  # a package, so that submodule imports go through the finder as well

        def __getattr__(self, item):
            if item.startswith("__"):
                raise AttributeError(item)
            return _Placeholder(f"{self.__name__}.{item}")

    class _Finder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
        def find_spec(self, fullname, path=None, target=None):
            if fullname == "triton_dist" or fullname.startswith("triton_dist."):
                return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
            return None

        def create_module(self, spec):
            return _StubModule(spec.name)

        def exec_module(self, module):
            pass

    assert "triton_dist" not in sys.modules, "the reference arm must not have this repository's package loaded"
    sys.meta_path.insert(0, _Finder())
    spec = importlib.util.spec_from_file_location("reference_tp_mlp", str(tree / "layers" / "nvidia" / "tp_mlp.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.TP_MLP


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


def decode_section(rank: int, world: int, steps: int, out: dict):
    """Qwen3-32B decode latency at TP=world: bs=1, ctx=512, megakernel (one launch per token). Results are written into
    `out` as they are measured, so that a time limit on the section keeps what was finished."""
    import torch
    import torch.distributed as dist

    from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
    from triton_dist.models import DenseLLM, ModelConfig
    from triton_dist.utils import perf_func_max_over_ranks

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
    out.update({"model": "Qwen3-32B (random init)", "tp": world, "bsz": 1, "ctx": 512, "ms_per_token": round(ms, 4),
                "weight_bytes_per_token_per_rank": wbytes, "hbm_floor_ms_of_" + pk["source"]: round(floor, 4),
                "fraction_of_hbm_floor": round(floor / ms, 3), "launches_per_token": 1, "graph": info,
                "reference_published_ms_8xH800": 7.41 if world == 8 else None})
    try:  # per-op-type timeline of one token (in-kernel task stamps), outside the timed region
        builder.enable_profiling(True)
        step()
        torch.cuda.synchronize()
        out["task_profile"] = builder.profile_summary()
        builder.enable_profiling(False)
    except Exception as e:  # noqa: BLE001
        out["task_profile"] = f"unavailable: {e}"
    mega.finalize()
    # same-box comparators: the other three columns of the reference's table (docs/mega_triton_kernel.md:29-33), through
    # Engine's CUDA-graph decode step on the SAME weights: torch (cuBLAS + NCCL), triton_dist_AR, triton_dist_gemm_ar
    if os.environ.get("TDIST_BENCH_ENGINE", "1") != "0":
        out["engine_graph_ms_per_token"] = {}
        engine_comparators(dense, cfg, steps, out["engine_graph_ms_per_token"])
        ref_ms = out["engine_graph_ms_per_token"].get("torch")
        if isinstance(ref_ms, float):
            out["megakernel_vs_torch_graph"] = round(ref_ms / ms, 3)
    del dense
    torch.cuda.empty_cache()
    return out


def engine_comparators(dense, cfg, steps: int, res: dict) -> dict:
    import torch

    from triton_dist.models import Engine
    from triton_dist.utils import perf_func_max_over_ranks

    ids = torch.randint(0, 32000, (1, 512), device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
    # torch + graph and the fused all-reduce backend were measured through this function at 1 and 8 GPUs. The fused
    # GEMM+AllReduce backend is opt-in here (TDIST_BENCH_GEMM_AR=1): its layer construction was corrected after the last
    # multi-GPU run, and an unverified path has no place inside the run that produces the headline line.
    backends = ("torch", "triton_dist_AR") + (("triton_dist_gemm_ar",) if os.environ.get("TDIST_BENCH_GEMM_AR") == "1" else ())
    for backend in backends:
        try:
            eng = Engine(cfg, model=dense)
            eng.backend = backend
            with torch.inference_mode():
                eng._init_kv_cache(bsz=1)
                eng.kv_cache.clear()
                dense.set_fwd(mode="torch")  # prefill on the torch path, as Engine.serve does
                dense.inference(input_ids=ids, position_ids=eng.get_ctx(ids), kv_cache=eng.kv_cache)
                eng.kv_cache.kv_offset.fill_(512)
                eng.set_backend(backend, 1)
                run = eng._init_cuda_graph(1)
                tok = torch.full((1, 1), 1225, dtype=torch.long, device="cuda")
                pos = torch.full((1, 1), 512, dtype=torch.long, device="cuda")
                _, ms = perf_func_max_over_ranks(lambda: run(tok, pos), steps, 3)
            res[backend] = round(ms, 4)
            eng.kv_cache = None
        except Exception as e:  # noqa: BLE001 - one comparator failing must not cost the others
            res[backend] = f"error: {type(e).__name__}: {e}"[:200]
        torch.cuda.empty_cache()
    dense.set_fwd(mode="torch")
    return res


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="tdist", choices=["tdist", "torch", "reference"])
    ap.add_argument("--no-decode", action="store_true", help="skip the Qwen3-32B decode section")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", str(1))
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29555")

    ref_mlp_cls = None
    if args.impl == "reference":
        tree = reference_tree()
        why = None
        if tree is None:
            why = "no reference source tree (baseline/_ref/triton_dist missing: run scripts/install_reference.sh)"
        else:
            try:
                ref_mlp_cls = load_reference_tp_mlp(tree)
            except Exception as e:  # noqa: BLE001
                why = f"reference layers/nvidia/tp_mlp.py failed to load: {type(e).__name__}: {e}"[:200]
        if why is not None:
            if int(os.environ["RANK"]) == 0:
                print(json.dumps({"impl": "reference", "unavailable": why}))
            return
    else:
        sys.path.insert(0, str(ROOT))

    import torch
    import torch.distributed as dist

    if args.impl == "reference":
        # plain torch.distributed: nothing of this repository runs in the reference arm
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
        finalize = dist.destroy_process_group
    else:
        from triton_dist.utils import finalize_distributed, initialize_distributed

        initialize_distributed(seed=1234)
        finalize = finalize_distributed
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
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)
    y_host = [torch.empty((m_loc, K_HID), dtype=dtype).pin_memory() for _ in range(2)]
    cleanup = []

    if args.impl == "tdist":
        from triton_dist.kernels.nvidia.allgather_gemm import ag_gemm, create_ag_gemm_context
        from triton_dist.kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs_op

        ag_ctx = create_ag_gemm_context(x_dev, w_up, rank, world, max_M=M, num_local_ranks=world)
        rs_ctx = create_gemm_rs_context(M, K_HID, rank, world, world, dtype)
        cleanup += [ag_ctx.finalize, rs_ctx.finalize]
        h = torch.empty((M, n_loc), dtype=dtype, device=dev)
        y = torch.empty((m_loc, K_HID), dtype=dtype, device=dev)

        def op_ag(x):
            return ag_gemm(x, w_up, ag_ctx, out=h)

        def op_rs(hh):
            return gemm_rs_op(hh, w_down, rs_ctx, output=y)
        launches_per_step = 2
    elif args.impl == "torch":
        from triton_dist.kernels.nvidia.allgather_gemm import torch_ag_gemm
        from triton_dist.kernels.nvidia.gemm_reduce_scatter import torch_gemm_rs

        def op_ag(x):
            return torch_ag_gemm(x, w_up)

        def op_rs(hh):
            return torch_gemm_rs(hh, w_down)
        launches_per_step = 0
    else:
        mlp = ref_mlp_cls(rank=rank, world_size=world, group=dist.group.WORLD)
        mlp.gate_up_proj, mlp.down_proj = w_up, w_down  # what TP_MLP._init_parameters leaves behind: the local shards
        op_ag, op_rs = mlp.torch_ag_gemm, mlp.torch_gemm_rs
        launches_per_step = 0

    def step(x):
        return op_rs(op_ag(x))

    # in-process comparator of the tdist arm: the reference's oracle path (NCCL + cuBLAS through torch)
    def baseline_step(x):
        from triton_dist.kernels.nvidia.allgather_gemm import torch_ag_gemm
        from triton_dist.kernels.nvidia.gemm_reduce_scatter import torch_gemm_rs

        return torch_gemm_rs(torch_ag_gemm(x, w_up), w_down)

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

    # end-to-end: every step copies its input from pinned host memory and reads the WHOLE result tensor back into pinned
    # host memory. Input copies run on one copy stream into two alternating device buffers, result copies on another
    # out of two alternating device buffers, so that H2D of step i+1 and D2H of step i-1 overlap the compute of step i
    # (what a serving loop does); the timed span covers all copies, all compute and all read-backs of the K steps.
    h2d_stream, d2h_stream = torch.cuda.Stream(), torch.cuda.Stream()
    x_bufs = [torch.empty_like(x_dev), torch.empty_like(x_dev)]
    y_bufs = [torch.empty((m_loc, K_HID), dtype=dtype, device=dev) for _ in range(2)]

    def timed_e2e(n_steps: int) -> float:
        cur = torch.cuda.current_stream()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        produced = [torch.cuda.Event(), torch.cuda.Event()]
        drained = [torch.cuda.Event(), torch.cuda.Event()]
        h2d_stream.wait_stream(cur)
        d2h_stream.wait_stream(cur)
        start.record(cur)
        h2d_stream.wait_event(start)
        d2h_stream.wait_event(start)
        for i in range(n_steps):
            b = i & 1
            with torch.cuda.stream(h2d_stream):
                if i >= 2:
                    h2d_stream.wait_event(consumed[b])  # buffer b is free once step i-2 has read it
                x_bufs[b].copy_(x_host, non_blocking=True)  # H2D of this step's input
                ready[b].record(h2d_stream)
            cur.wait_event(ready[b])
            if i >= 2:
                cur.wait_event(drained[b])  # result buffer b has been read back
            out = step(x_bufs[b])
            y_bufs[b].copy_(out)  # device-side hand-over (the op's output buffer is reused by the next step)
            consumed[b].record(cur)
            produced[b].record(cur)
            with torch.cuda.stream(d2h_stream):
                d2h_stream.wait_event(produced[b])
                y_host[b].copy_(y_bufs[b], non_blocking=True)  # D2H of the step's result [M/W, 4096]
                drained[b].record(d2h_stream)
        cur.wait_stream(d2h_stream)
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
    ag_ms = op_ms(lambda: op_ag(x_dev))
    h_in = op_ag(x_dev)
    rs_ms = op_ms(lambda: op_rs(h_in))
    base_ms = op_ms(lambda: baseline_step(x_dev)) if args.impl == "tdist" else None

    flops_step = 2.0 * M * N_FFN * K_HID * 2  # whole job, both GEMMs
    ms_per_step = total_ms / args.steps
    value = flops_step / (ms_per_step * 1e-3) / 1e12
    pk = measured_peaks()
    per_gpu_flops = flops_step / 2 / world
    link_bytes = (world - 1) * m_loc * K_HID * 2
    roof_ms = max(per_gpu_flops / (pk["bf16_tflops"] * 1e9), link_bytes / (NVLINK_GBS * 1e6))
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
                "ms_per_step": round(e2e_ms / args.steps, 5), "h2d_bytes_per_step": x_host.numel() * 2,
                "d2h_bytes_per_step": y_host[0].numel() * 2,
                "note": "per rank; H2D of step i+1 and D2H of step i-1 overlap the compute of step i (two copy streams, "
                        "double-buffered device and pinned host buffers); the whole result [M/W, 4096] is read back"},
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
    if args.impl == "reference":
        out["reference_class"] = REFERENCE_CLASS
        out["reference_source"] = str(reference_tree())
    for fn in cleanup:
        fn()
    if args.impl == "tdist" and not args.no_decode and os.environ.get("TDIST_BENCH_DECODE", "1") != "0":
        # The headline number must survive the secondary section, whatever happens in it: an exception is recorded, and a
        # section that does not come back within the limit (every rank runs the same timer) ends the process after rank 0
        # has printed the line without it.
        # (600 s: the round-1 driver gave every N of its scaling run 870 s in total, and the headline part needs about a minute)
        limit_s = float(os.environ.get("TDIST_BENCH_DECODE_TIMEOUT_S", "600"))

        partial: dict = {}

        def give_up():
            if rank == 0:
                try:
                    out["decode"] = dict(partial, error=f"decode section did not finish within {limit_s:.0f} s; what was measured until then is kept")
                    line = json.dumps(out)
                except Exception:  # noqa: BLE001 - (the section was writing into `partial` at this very moment)
                    out["decode"] = {"error": f"decode section did not finish within {limit_s:.0f} s"}
                    line = json.dumps(out)
                print(line, flush=True)
            os._exit(0)

        timer = threading.Timer(limit_s, give_up)
        timer.daemon = True
        timer.start()
        try:
            out["decode"] = decode_section(rank, world, max(args.steps, 10), partial)
        except Exception as e:
            out["decode"] = dict(partial, error=f"{type(e).__name__}: {e}"[:300])
        timer.cancel()
    if rank == 0:
        print(json.dumps(out), flush=True)
    try:
        finalize()
    except Exception as e:  # tearing down after a failed secondary section must not turn a printed result into a failed run
        print(f"bench: finalize failed: {type(e).__name__}: {e}", file=sys.stderr, flush=True)
        os._exit(0)


if __name__ == "__main__":
    main()
