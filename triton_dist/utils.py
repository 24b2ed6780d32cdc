"""Host runtime helpers — same public surface as the reference's
``python/triton_dist/utils.py``, re-implemented on top of the
native symmetric heap instead of NVSHMEM.

The ``nvshmem_*`` names are kept as aliases so that code written against the
reference keeps working; none of them touches NVSHMEM here.
"""
from __future__ import annotations

import contextlib
import datetime
import functools
import gzip
import json
import os
import random
import subprocess
import sys
import time
from pathlib import Path
from typing import Callable, Iterable, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .runtime import symm as _symm

NVSHMEM_SIGNAL_DTYPE = torch.int32  # flags are 32-bit in this framework (reference: int64, utils.py:941)
SIGNAL_DTYPE = NVSHMEM_SIGNAL_DTYPE

_TP_GROUP: Optional[dist.ProcessGroup] = None
_CPU_GROUP: Optional[dist.ProcessGroup] = None


# --------------------------------------------------------------------------- platform probes
def is_cuda() -> bool:
    return torch.cuda.is_available() and torch.version.hip is None


def is_hip() -> bool:
    return False  # B200-only framework


def has_tma() -> bool:
    return is_cuda() and torch.cuda.get_device_capability()[0] >= 9


def is_blackwell() -> bool:
    return is_cuda() and torch.cuda.get_device_capability()[0] >= 10


def get_device_property(device_id: int = 0):
    return torch.cuda.get_device_properties(device_id)


def get_device_max_shared_memory_size(device: int = 0) -> int:
    props = torch.cuda.get_device_properties(device)
    return int(getattr(props, "shared_memory_per_block_optin", 227 * 1024))


def requires(condition_func: Callable[[], bool]):
    """Decorator: raise at call time if a platform capability is missing."""

    def decorator(func):
        @functools.wraps(func)
        def wrapper(*args, **kwargs):
            if not condition_func():
                raise RuntimeError(f"{func.__name__} requires {condition_func.__name__}() to hold on this platform")
            return func(*args, **kwargs)

        return wrapper

    return decorator


def launch_cooperative_grid_options() -> dict:
    """Kept for API compatibility: native launches set the cooperative attribute themselves."""
    return {}


# --------------------------------------------------------------------------- bootstrap
def init_seed(seed: int = 0) -> None:
    """Deterministic seeds, TF32 off (reference utils.py:77-96)."""
    os.environ.setdefault("NCCL_DEBUG", os.getenv("NCCL_DEBUG", "ERROR"))
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    np.random.seed(seed)
    random.seed(seed)


def _env_int(name: str, default: int) -> int:
    return int(os.environ.get(name, default))


def bind_to_gpu_numa_node(local_rank: int) -> bool:
    """Pin this process to the CPU cores next to its GPU (NVML's ideal affinity), so that pinned host buffers are
    first-touched on the GPU's own NUMA node: with one process per GPU on a two-socket host, unpinned ranks may share
    one socket's memory and PCIe root, and the host<->device copies of all ranks slow each other down. Opt-in with
    TDIST_NUMA_BIND=1 (never narrows the process to fewer than 8 cores). The reference leaves placement to the
    launcher (scripts/launch.sh)."""
    if os.environ.get("TDIST_NUMA_BIND", "0") != "1":
        return False
    try:
        import pynvml

        pynvml.nvmlInit()
        try:  # CUDA_VISIBLE_DEVICES may renumber the devices: find the NVML handle by UUID
            h = pynvml.nvmlDeviceGetHandleByUUID(f"GPU-{torch.cuda.get_device_properties(local_rank).uuid}")
        except Exception:  # noqa: BLE001
            h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if len(cpus) < 8:
            return False
        os.sched_setaffinity(0, cpus)
        return True
    except Exception:  # noqa: BLE001 - placement is an optimisation, never a requirement
        return False


def initialize_distributed(seed: Optional[int] = None, heap_bytes: Optional[int] = None) -> dist.ProcessGroup:
    """torch.distributed + symmetric-heap bring-up (reference utils.py:182-205).

    One process per GPU. NCCL is used for glue collectives and as the baseline,
    gloo for CPU-side object exchange; the heap replaces NVSHMEM's init-by-UID.
    Without a GPU the gloo-only plumbing backend is initialised instead.
    """
    global _TP_GROUP, _CPU_GROUP
    rank, world, local_rank = _env_int("RANK", 0), _env_int("WORLD_SIZE", 1), _env_int("LOCAL_RANK", 0)
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(local_rank)
        bind_to_gpu_numa_node(local_rank)
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        kwargs = dict(backend="cpu:gloo,cuda:nccl" if cuda else "gloo", world_size=world, rank=rank,
                      timeout=datetime.timedelta(seconds=1800))
        if cuda:
            kwargs["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(**kwargs)
    _TP_GROUP = dist.new_group(ranks=list(range(world)), backend="cpu:gloo,cuda:nccl" if cuda else "gloo")
    _CPU_GROUP = dist.new_group(ranks=list(range(world)), backend="gloo")
    dist.barrier(group=_CPU_GROUP)
    init_seed(rank if seed is None else seed)
    _symm.init_heap(rank, world, local_rank, _CPU_GROUP, heap_bytes)
    return _TP_GROUP


def init_nvshmem_by_torch_process_group(pg: dist.ProcessGroup) -> None:
    """API-compat alias: brings up the symmetric heap over ``pg`` (gloo-capable)."""
    rank, world, local_rank = dist.get_rank(pg), dist.get_world_size(pg), _env_int("LOCAL_RANK", 0)
    _symm.init_heap(rank, world, local_rank, pg)


def finalize_distributed() -> None:
    _symm.finalize_heap()
    if dist.is_initialized():
        dist.destroy_process_group()


def get_tp_group() -> Optional[dist.ProcessGroup]:
    return _TP_GROUP


def get_cpu_group() -> Optional[dist.ProcessGroup]:
    return _CPU_GROUP


# --------------------------------------------------------------------------- symmetric tensors
def symm_tensor(shape: Sequence[int], dtype: torch.dtype, zero: bool = True) -> torch.Tensor:
    """Allocate a symmetric tensor (collective: same call order on all ranks)."""
    return _symm.get_heap().alloc(tuple(shape) if not isinstance(shape, int) else (shape,), dtype, zero)


def nvshmem_create_tensor(shape, dtype) -> torch.Tensor:
    return symm_tensor(shape, dtype)


def nvshmem_create_tensors(shape, dtype, rank: int, local_world_size: int) -> List[torch.Tensor]:
    """List of views of one symmetric tensor, index = peer local rank (utils.py:121-136)."""
    heap = _symm.get_heap()
    t = symm_tensor(shape, dtype)
    if isinstance(heap, _symm.HostHeap):
        return [t for _ in range(local_world_size)]
    node_base = rank // local_world_size * local_world_size
    return [heap.peer_view(t, node_base + p) for p in range(local_world_size)]


def nvshmem_free_tensor_sync(tensor: Union[torch.Tensor, Sequence[torch.Tensor]]) -> None:
    heap = _symm.get_heap()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    if isinstance(tensor, (list, tuple)):
        tensor = tensor[heap.rank % len(tensor)] if len(tensor) else None
    if tensor is not None:
        heap.free(tensor)


def symm_peer_view(t: torch.Tensor, peer: int) -> torch.Tensor:
    return _symm.get_heap().peer_view(t, peer)


class TorchStreamWrapper:
    """Exposes a torch stream through the ``__cuda_stream__`` protocol (utils.py:151-160)."""

    def __init__(self, pt_stream: "torch.cuda.Stream"):
        self.pt_stream = pt_stream
        self.handle = pt_stream.cuda_stream

    def __cuda_stream__(self):
        return (0, self.handle)


def nvshmem_barrier_all_on_stream(stream: Optional["torch.cuda.Stream"] = None) -> None:
    """Stream-ordered barrier over all ranks; unlike NVSHMEM's host barrier it is graph-capturable."""
    _symm.get_heap().barrier_all(stream)


barrier_all_on_stream = nvshmem_barrier_all_on_stream


def nvshmem_signal_wait(signal: torch.Tensor, pe: int, signal_val: int, signal_op: int,
                        stream: Optional["torch.cuda.Stream"] = None) -> None:
    """Set ``signal`` on rank ``pe`` and wait for the same value locally (reference utils.py:170-179)."""
    from .kernels.common_ops import set_signal, wait_eq

    heap = _symm.get_heap()
    set_signal(heap.peer_view(signal, pe), signal_val, stream, add=(signal_op == 10))
    wait_eq(signal, signal_val, stream)


# --------------------------------------------------------------------------- data / checks
@contextlib.contextmanager
def with_torch_deterministic(mode: bool, warn_only: bool = True):
    old = torch.are_deterministic_algorithms_enabled()
    torch.use_deterministic_algorithms(mode, warn_only=warn_only)
    try:
        yield
    finally:
        torch.use_deterministic_algorithms(old, warn_only=warn_only)


def is_fp8_dtype(dtype: torch.dtype) -> bool:
    return dtype.itemsize == 1 and dtype.is_floating_point


def default_device() -> str:
    return "cuda" if torch.cuda.is_available() else "cpu"


def _make_tensor(shape, dtype: torch.dtype, init_args: Tuple[float, float], device: Optional[str] = None):
    device = device or default_device()
    if callable(shape):
        shape = shape()
    if isinstance(shape, int):
        shape = (shape,)
    shape = tuple(s() if callable(s) else s for s in shape)
    scale, bias = init_args
    if dtype in (torch.float16, torch.bfloat16, torch.float32):
        return (torch.rand(shape, dtype=dtype, device=device) * 2 - 1) * scale + bias
    if dtype == torch.int8:
        return torch.randint(-int(scale), int(scale), shape, dtype=torch.int8, device=device) + int(bias)
    if is_fp8_dtype(dtype):
        return ((torch.rand(shape, dtype=torch.float16, device=device) * 2 - 1) * scale + bias).to(dtype)
    raise ValueError(f"unsupported dtype {dtype}")


def generate_data(configs):
    """Infinite generator of fresh random tensors, one per ``(shape, dtype, (scale, bias))`` entry."""
    while True:
        yield tuple(_make_tensor(*args) if args else None for args in configs)


def assert_allclose(x: torch.Tensor, y: torch.Tensor, rtol: float, atol: float, verbose: bool = True) -> None:
    if torch.allclose(x, y, rtol=rtol, atol=atol):
        if verbose:
            print("all close")
        return
    bad = ~torch.isclose(x, y, rtol=rtol, atol=atol)
    n_bad = int(bad.sum())
    diff = (x.float() - y.float()).abs()
    idx = bad.nonzero(as_tuple=False)[:8].tolist()
    msg = (f"tensors differ: {n_bad}/{y.numel()} ({100.0 * n_bad / max(y.numel(), 1):.3f}%) elements, "
           f"max |x-y| = {float(diff.max()):.6g} (atol {atol}, rtol {rtol}), shape {list(y.shape)}, first bad at {idx}")
    print(msg, file=sys.stderr)
    raise RuntimeError(msg)


def bitwise_equal(x: torch.Tensor, y: torch.Tensor) -> bool:
    return x.shape == y.shape and bool(torch.equal(x.contiguous().view(torch.uint8), y.contiguous().view(torch.uint8)))


# --------------------------------------------------------------------------- timing / printing
def sleep_async(duration_ms: float) -> None:
    """Queue a busy-wait on the current stream (hides launch latency before timing)."""
    if torch.cuda.is_available():
        hz = torch.cuda.clock_rate() * 1e6 if hasattr(torch.cuda, "clock_rate") else 1.9e9
        torch.cuda._sleep(int(hz * duration_ms / 1000))


def perf_func(func: Callable, iters: int, warmup_iters: int):
    """CUDA-event timing of ``iters`` back-to-back calls; returns (last output, ms per call)."""
    if not torch.cuda.is_available():
        out = None
        for _ in range(warmup_iters):
            out = func()
        t0 = time.perf_counter()
        for _ in range(iters):
            out = func()
        return out, (time.perf_counter() - t0) * 1e3 / max(iters, 1)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = None
    for n in range(iters + warmup_iters):
        if n == warmup_iters:
            start.record()
        out = func()
    stop.record()
    torch.cuda.current_stream().synchronize()
    return out, start.elapsed_time(stop) / max(iters, 1)


def perf_func_max_over_ranks(func: Callable, iters: int, warmup_iters: int, group=None, flush_l2: bool = False,
                             use_graph: bool = False):
    """Device-timed ms per call, MAX over ranks (what every multi-GPU number in this repo reports).

    ``flush_l2``: a 256 MiB write evicts L2 before every call; each call is then timed by its own event
    pair so that the flush is outside the measured span. ``use_graph``: capture ``iters`` calls in one CUDA
    graph (removes host launch overhead; for latency-bound ops)."""
    if dist.is_initialized():
        dist.barrier(group=group)
    out = None
    if not torch.cuda.is_available():
        out, ms = perf_func(func, iters, warmup_iters)
    elif use_graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(warmup_iters, 3)):
                out = func()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(iters):
                out = func()
        g.replay()
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier(group=group)
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        g.replay()
        stop.record()
        torch.cuda.synchronize()
        ms = start.elapsed_time(stop) / iters
    elif flush_l2:
        flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")
        for _ in range(warmup_iters):
            out = func()
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier(group=group)
        evs = []
        for _ in range(iters):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            out = func()
            e.record()
            evs.append((s, e))
        torch.cuda.synchronize()
        ms = sum(s.elapsed_time(e) for s, e in evs) / max(iters, 1)
    else:
        torch.cuda.synchronize()
        out, ms = perf_func(func, iters, warmup_iters)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=default_device())
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        ms = float(t.item())
    return out, ms


def dist_print(*args, allowed_ranks: Union[str, Iterable[int]] = (0,), prefix: bool = False, need_sync: bool = False,
               **kwargs) -> None:
    rank, world = _env_int("RANK", 0), _env_int("WORLD_SIZE", 1)
    ranks = range(world) if allowed_ranks == "all" else allowed_ranks
    for r in ranks:
        if need_sync and dist.is_initialized():
            dist.barrier()
        if rank == r:
            if prefix:
                print(f"[rank:{rank}]", end="")
            print(*args, **kwargs)
            sys.stdout.flush()


def CUDA_CHECK(err) -> None:
    """Accepts cuda-python (err, ...) tuples / enums or plain ints."""
    if isinstance(err, tuple):
        err = err[0]
    code = int(getattr(err, "value", err))
    if code != 0:
        raise RuntimeError(f"CUDA error: {err}")


# --------------------------------------------------------------------------- profiling
_NVTX = os.environ.get("TDIST_NVTX", "0") == "1"


def launch_metadata(name: str, describe: Optional[Callable[..., dict]] = None):
    """Name the launches of an op for profilers, the way the reference's kernels carry ``launch_metadata`` (kernel name with
    M / N / K, flops, bytes: allgather_gemm.py:145-155) for Proton. With ``TDIST_NVTX=1`` every call of the decorated op is
    wrapped in an NVTX range ``name[k=v,...]`` (visible in ncu / Nsight Systems next to the kernels it launched);
    ``describe(*args, **kwargs)`` returns the fields (shapes, ``flops``, ``bytes``). Without the variable the decorator
    returns the function itself: no wrapper, no cost."""

    def deco(fn):
        if not _NVTX:
            return fn

        @functools.wraps(fn)
        def wrapper(*args, **kwargs):
            if not torch.cuda.is_available():
                return fn(*args, **kwargs)
            label = name
            if describe is not None:
                try:
                    label = f"{name}[{','.join(f'{k}={v}' for k, v in describe(*args, **kwargs).items())}]"
                except Exception:  # noqa: BLE001 - a label must never break the op
                    pass
            torch.cuda.nvtx.range_push(label)
            try:
                return fn(*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()

        return wrapper

    return deco


def gemm_metadata(M: int, N: int, K: int, elt: int = 2, world: int = 1, comm_bytes: int = 0) -> dict:
    """Fields of a GEMM-bearing op: per-rank flops, HBM bytes of the operands and the bytes that cross NVLink."""
    return {"M": M, "N": N, "K": K, "W": world, "flops": 2 * M * N * K, "bytes": (M * K + N * K + M * N) * elt, "link_bytes": comm_bytes}


def get_torch_prof_ctx(do_prof: bool):
    if not do_prof:
        return contextlib.nullcontext()
    acts = [torch.profiler.ProfilerActivity.CPU]
    if torch.cuda.is_available():
        acts.append(torch.profiler.ProfilerActivity.CUDA)
    return torch.profiler.profile(activities=acts, record_shapes=True, with_stack=False)


class group_profile:
    """torch.profiler around a block, one chrome trace per rank, merged on rank 0
    into ``prof/<name>/merged.json.gz`` with one pid lane per rank
    (reference utils.py:505-589)."""

    def __init__(self, name: Optional[str] = None, do_prof: bool = True, merge_group: bool = True, keep_merged_only: bool = True,
                 group: Optional[dist.ProcessGroup] = None, out_dir: str = "prof"):
        self.name = name or "trace"
        self.do_prof = do_prof
        self.merge_group = merge_group
        self.keep_merged_only = keep_merged_only
        self.group = group
        self.out_dir = Path(out_dir) / self.name
        self.prof = None

    def __enter__(self):
        if self.do_prof:
            self.prof = get_torch_prof_ctx(True)
            self.prof.__enter__()
        return self

    def __exit__(self, exc_type, exc, tb):
        if not self.do_prof:
            return False
        self.prof.__exit__(exc_type, exc, tb)
        rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.out_dir.mkdir(parents=True, exist_ok=True)
        mine = self.out_dir / f"rank{rank}.json"
        self.prof.export_chrome_trace(str(mine))
        if self.merge_group and world > 1:
            self._merge(rank, world, mine)
        return False

    def _merge(self, rank: int, world: int, mine: Path) -> None:
        events = json.loads(mine.read_text()).get("traceEvents", [])
        for ev in events:  # one process lane per rank
            if isinstance(ev.get("pid"), int):
                ev["pid"] = ev["pid"] + rank * 10_000_000
            elif "pid" in ev:
                ev["pid"] = f"rank{rank}:{ev['pid']}"
        gathered = [None] * world if rank == 0 else None
        dist.gather_object(events, gathered, dst=0, group=self.group)
        if rank == 0:
            merged = {"traceEvents": [e for evs in gathered for e in evs]}
            with gzip.open(self.out_dir / "merged.json.gz", "wt") as f:
                json.dump(merged, f)
        if self.keep_merged_only:
            mine.unlink(missing_ok=True)


# --------------------------------------------------------------------------- topology probes (nvidia-smi / NVML)
def _nvml():
    try:
        import pynvml

        pynvml.nvmlInit()
        return pynvml
    except Exception:
        return None


@functools.lru_cache()
def has_fullmesh_nvlink() -> bool:
    """True when every GPU pair is connected by NVLink (always the case behind NVSwitch)."""
    if not torch.cuda.is_available():
        return False
    n = torch.cuda.device_count()
    if n <= 1:
        return True
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        rows = [ln.split() for ln in out.splitlines() if ln.startswith("GPU")]
        for i, row in enumerate(rows[:n]):
            for j, cell in enumerate(row[1:1 + n]):
                if i != j and not cell.startswith("NV"):
                    return False
        return bool(rows)
    except Exception:
        return all(torch.cuda.can_device_access_peer(0, j) for j in range(1, n))


@functools.lru_cache()
def get_nvlink_max_speed(gpu_index: int = 0) -> float:
    """Aggregate NVLink bandwidth per direction in GB/s (NVML; nominal 900 for NVLink 5 x18)."""
    nv = _nvml()
    if nv is not None:
        try:
            h = nv.nvmlDeviceGetHandleByIndex(gpu_index)
            total = 0.0
            for link in range(18):
                try:
                    if nv.nvmlDeviceGetNvLinkState(h, link):
                        fv = nv.nvmlDeviceGetFieldValues(h, [(nv.NVML_FI_DEV_NVLINK_SPEED_MBPS_L0 + link)])[0]
                        total += float(fv.value.uiVal) / 1000.0
                except Exception:
                    break
            if total > 0:
                return total
        except Exception:
            pass
    return 900.0


def get_intranode_max_speed(gpu_index: int = 0, with_scale: bool = False) -> float:
    bw = get_nvlink_max_speed(gpu_index) if has_fullmesh_nvlink() else 64.0
    return bw * (0.85 if with_scale else 1.0)


def get_numa_world_size() -> int:
    """NVSwitch is flat: the whole node is one NUMA group for the purpose of collective hierarchy."""
    return _env_int("LOCAL_WORLD_SIZE", torch.cuda.device_count() if torch.cuda.is_available() else 1)


@functools.lru_cache()
def supports_p2p_native_atomic() -> bool:
    if not torch.cuda.is_available():
        return False
    n = torch.cuda.device_count()
    return all(torch.cuda.can_device_access_peer(0, j) for j in range(1, n))


def is_nvshmem_multimem_supported() -> bool:
    """NVLS multicast bound on the live heap (or supported by the device when no heap exists yet)."""
    if _symm.heap_initialized():
        return bool(_symm.get_heap().multicast)
    return is_blackwell() or (is_cuda() and torch.cuda.get_device_capability()[0] >= 9)


is_multimem_supported = is_nvshmem_multimem_supported


def measured_peaks() -> dict:
    """Driver-measured roofline denominators (MEASURED_PEAKS.json) or the profiling recipe's fallback."""
    p = Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


NVLINK_GBS_MEASURED = 770.0  # per direction per GPU, B200_PROFILING.md


def all_to_all_compat(outs: List[torch.Tensor], ins: List[torch.Tensor], group=None) -> None:
    """dist.all_to_all that also works on gloo (CPU plumbing tests): variable-size exchange through
    all_gather_object when the backend has no alltoall."""
    if ins[0].is_cuda:
        dist.all_to_all(outs, ins, group=group)
        return
    rank = dist.get_rank(group)
    world = dist.get_world_size(group)
    gathered: List[Optional[list]] = [None] * world
    dist.all_gather_object(gathered, [t.clone() for t in ins], group=group)
    for s in range(world):
        outs[s].copy_(gathered[s][rank].reshape(outs[s].shape))
