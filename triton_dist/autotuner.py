"""Contextual autotuner (reference ``python/triton_dist/autotuner.py:43-250``).

The reference monkey-patches Triton's ``Autotuner.run`` so that a *whole thunk* (possibly several kernels plus
communication) is re-run until every inner tunable has tried all its configs, and all-reduces the timings with MAX over
ranks so that every rank picks the same config (``autotuner.py:220-231``). Native kernels are ahead-of-time compiled
template instances, so a "config" here is a set of launch knobs (tile width, number of communication CTAs, split
count, method). Usage:

    @tunable("ag_gemm", {"BLOCK_N": [0, 224, 256], "num_comm_sms": [12, 16, 20, 24]})
    def my_ag_gemm(a, b, ctx, BLOCK_N=0, num_comm_sms=-1): ...

    best = contextual_autotune(is_dist=True)(lambda: my_ag_gemm(a, b, ctx))()   # runs the sweep, caches the winner

Cross-rank consensus: the per-config time is the MAX over ranks; ties are broken by config order, so every rank selects
the same winner. Results are cached per (name, key) and logged to ``./.autotune_logs/rank-N.log`` like the reference.
"""
from __future__ import annotations

import functools
import itertools
import os
import time
from pathlib import Path
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

_REGISTRY: Dict[str, "Tunable"] = {}
_ACTIVE: Optional["_Session"] = None


def _log(msg: str) -> None:
    d = Path(".autotune_logs")
    try:
        d.mkdir(exist_ok=True)
        with open(d / f"rank-{os.environ.get('RANK', '0')}.log", "a") as f:
            f.write(msg + "\n")
    except OSError:
        pass


class Tunable:
    """A function with a discrete configuration space; picks its config from the active session or its cache."""

    def __init__(self, name: str, space: Dict[str, List[Any]], fn: Callable, key: Optional[Callable[..., Any]] = None):
        self.name, self.fn, self.key_fn = name, fn, key
        self.configs = [dict(zip(space.keys(), vals)) for vals in itertools.product(*space.values())]
        self.best: Dict[Any, Dict[str, Any]] = {}
        functools.update_wrapper(self, fn)

    def _key(self, args, kwargs):
        if self.key_fn is not None:
            return self.key_fn(*args, **kwargs)
        return tuple(tuple(a.shape) + (str(a.dtype),) for a in args if isinstance(a, torch.Tensor))

    def __call__(self, *args, **kwargs):
        key = self._key(args, kwargs)
        if _ACTIVE is not None:
            cfg = _ACTIVE.config_for(self, key)
        else:
            cfg = self.best.get(key, {})
        return self.fn(*args, **{**kwargs, **cfg})


def tunable(name: str, space: Dict[str, List[Any]], key: Optional[Callable[..., Any]] = None):
    def deco(fn):
        t = Tunable(name, space, fn, key)
        _REGISTRY[name] = t
        return t

    return deco


class _Session:
    """One sweep: every tunable reached by the thunk walks through its configs in lock-step (mixed-radix counter)."""

    def __init__(self):
        self.seen: List[Tuple[Tunable, Any]] = []
        self.index: Dict[Tuple[int, Any], int] = {}

    def config_for(self, t: Tunable, key) -> Dict[str, Any]:
        k = (id(t), key)
        if k not in self.index:
            self.index[k] = 0
            self.seen.append((t, key))
        return t.configs[self.index[k]]

    def advance(self) -> bool:
        """Next combination; False when the space is exhausted."""
        for t, key in self.seen:
            k = (id(t), key)
            self.index[k] += 1
            if self.index[k] < len(t.configs):
                return True
            self.index[k] = 0
        return False

    def snapshot(self) -> Dict[Tuple[int, Any], int]:
        return dict(self.index)


def _time_thunk(thunk: Callable, warmup: int, iters: int, is_dist: bool, group) -> float:
    if torch.cuda.is_available():
        for _ in range(warmup):
            thunk()
        if is_dist and dist.is_initialized():
            dist.barrier(group=group)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            thunk()
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / iters
    else:
        for _ in range(warmup):
            thunk()
        t0 = time.perf_counter()
        for _ in range(iters):
            thunk()
        ms = (time.perf_counter() - t0) * 1e3 / iters
    if is_dist and dist.is_initialized() and dist.get_world_size(group) > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda" if torch.cuda.is_available() else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)  # consensus: the slowest rank defines the time
        ms = float(t.item())
    return ms


def contextual_autotune(is_dist: bool = False, warmup: int = 3, iters: int = 10, group=None, n_repeat: Optional[int] = None,
                        n_warmup: Optional[int] = None):
    """Decorator: tune every ``@tunable`` reached by the wrapped thunk, then run it once with the winners.
    ``n_repeat`` / ``n_warmup`` are the reference's names for ``iters`` / ``warmup`` (autotuner.py:97-102)."""
    if n_repeat is not None:
        iters = n_repeat
    if n_warmup is not None:
        warmup = n_warmup

    def deco(thunk: Callable):
        @functools.wraps(thunk)
        def wrapper(*args, **kwargs):
            global _ACTIVE
            sess = _Session()
            _ACTIVE = sess
            best_ms, best_snap = float("inf"), None
            try:
                while True:
                    ms = _time_thunk(lambda: thunk(*args, **kwargs), warmup, iters, is_dist, group)
                    snap = sess.snapshot()
                    desc = {t.name: t.configs[snap[(id(t), key)]] for t, key in sess.seen}
                    _log(f"{desc} -> {ms:.4f} ms")
                    if ms < best_ms:
                        best_ms, best_snap = ms, snap
                    if not sess.advance():
                        break
            finally:
                _ACTIVE = None
            for t, key in sess.seen:
                t.best[key] = t.configs[best_snap[(id(t), key)]]
                _log(f"best {t.name} {key}: {t.best[key]} ({best_ms:.4f} ms)")
            return thunk(*args, **kwargs)

        return wrapper

    return deco
