"""Shared harness for the script-style tests (run under torchrun, like the
reference's ``python/triton_dist/test/nvidia/*.py`` under ``scripts/launch.sh``)."""
from __future__ import annotations

import argparse
from typing import Optional
import os
import sys

import torch
import torch.distributed as dist

from triton_dist.utils import (dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks)

DTYPES = {"bfloat16": torch.bfloat16, "float16": torch.float16, "float32": torch.float32, "bf16": torch.bfloat16,
          "fp16": torch.float16, "fp32": torch.float32}

# tolerances of the reference's tests (test_gemm_rs.py:108-113, test_allreduce.py:99-103)
TOL = {torch.bfloat16: dict(atol=6e-2, rtol=6e-2), torch.float16: dict(atol=1e-2, rtol=1e-2),
       torch.float32: dict(atol=1e-3, rtol=1e-3)}


def base_parser(desc: str) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description=desc)
    p.add_argument("--dtype", default="bfloat16", choices=list(DTYPES))
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--iters", type=int, default=20)
    p.add_argument("--check", action="store_true", help="correctness only")
    p.add_argument("--verify-iters", type=int, default=4)
    p.add_argument("--stress", action="store_true", help="random skew between ranks every iteration")
    p.add_argument("--profile", action="store_true")
    p.add_argument("--graph", action="store_true", help="time inside a CUDA graph (no host launch overhead)")
    return p


_SUITE = {"active": False, "group": None, "results": []}


def setup(seed: int = 3407):
    if _SUITE["active"] and _SUITE["group"] is not None:  # suite mode: one process group / heap for many cases
        group = _SUITE["group"]
    else:
        group = initialize_distributed(seed=None)
        _SUITE["group"] = group
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(seed + rank)
    from triton_dist.runtime import get_heap
    h = get_heap()
    dist_print(f"[setup] world={world} heap_mode={h.mode} (1=VMM 2=cudaIpc -1=host) multicast={h.multicast}")
    return group, rank, world


def device() -> str:
    return "cuda" if torch.cuda.is_available() else "cpu"


def rand(shape, dtype, scale=1.0):
    return ((torch.rand(shape, dtype=torch.float32, device=device()) * 2 - 1) * scale).to(dtype)


def all_ok(ok: bool) -> bool:
    """Cross-rank AND of a pass/fail flag, so that every rank leaves a verify loop together (a rank that
    stopped calling a fused op would otherwise hang its peers inside the kernel)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device())
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok = bool(t.item())
    return ok


def check_close(x: torch.Tensor, ref_fp32: torch.Tensor, dtype, what: str = "") -> bool:
    """Compare against an fp32 ground truth with a tolerance scaled to the output magnitude: a few units of
    the output dtype's epsilon times max|ref| (partial sums are rounded to 16 bit before the cross-rank reduction,
    exactly as in the reference's symmetric-buffer design). Prints and returns False on mismatch."""
    eps = {torch.bfloat16: 2.0 ** -8, torch.float16: 2.0 ** -11, torch.float32: 2.0 ** -20}[dtype]
    scale = float(ref_fp32.abs().max().clamp_min(1e-6))
    err = (x.float() - ref_fp32).abs()
    tol = 6 * eps * scale
    bad = int((err > tol).sum())
    if bad:
        dist_print(f"{what}: {bad}/{x.numel()} elements off by more than {tol:.4g} (max err {float(err.max()):.4g}, "
                   f"max|ref| {scale:.4g})", allowed_ranks="all", prefix=True)
    return bad == 0


def logits_close(lg: torch.Tensor, ref: torch.Tensor, dtype, what: str = "", frac: Optional[float] = None) -> bool:
    """Model-level check against a reference implementation's logits [rows, vocab]. Probabilities of a random-init model
    are all ~1/vocab, so comparing softmax outputs with an absolute tolerance can hardly fail; instead
      * every logit must agree within a tolerance that is a fraction of the SPREAD of the reference logits of its row
        (10 % for 16-bit runs: activations are rounded to 16 bit in every layer; 0.2 % in fp32), and
      * the greedy token must be the same wherever the reference's top-1 leads its runner-up by more than that tolerance.
    A wrong weight shard, a missing all-reduce or a stale KV row moves logits by about one spread and fails both.
    ``frac`` overrides the fraction (16-bit MoE models: a rounding difference in the router can flip one expert of one
    token, which moves that token's logits by a few tenths of the spread without anything being wrong)."""
    lg, ref = lg.float(), ref.float()
    spread = ref.std(dim=-1, keepdim=True).clamp_min(1e-6)
    if frac is None:
        frac = 0.10 if dtype in (torch.bfloat16, torch.float16) else 2e-3
    tol = frac * spread
    err = (lg - ref).abs()
    ok_val = bool((err <= tol).all())
    top2 = ref.topk(2, dim=-1).values
    decided = (top2[:, 0] - top2[:, 1]) > 2 * tol[:, 0]
    ok_top = bool((lg.argmax(-1)[decided] == ref.argmax(-1)[decided]).all())
    if not (ok_val and ok_top):
        dist_print(f"{what}: max |logit err| / spread = {float((err / spread).max()):.4g} (limit {float((tol / spread).max()):.3g}), "
                   f"top-1 agreement on decided rows: {ok_top} ({int(decided.sum())}/{decided.numel()} rows decided)",
                   allowed_ranks="all", prefix=True)
    return ok_val and ok_top


def skew(enabled: bool) -> None:
    """Race/hang provocation: delay this rank by a random amount (reference allgather.py:74-78)."""
    if enabled and torch.cuda.is_available():
        torch.cuda._sleep(int(torch.randint(0, 3_000_000, (1,)).item()))


def finish(ok: bool = True):
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    dist.barrier()
    if _SUITE["active"]:
        _SUITE["results"].append(bool(ok))
        raise SystemExit(0 if ok else 1)
    dist_print("PASS" if ok else "FAIL")
    finalize_distributed()
    sys.exit(0 if ok else 1)


def run_suite(cases) -> bool:
    """Run several script-style tests in ONE process group (saves the ~20 s start-up per torchrun launch).
    ``cases``: list of (module, [argv...])."""
    import runpy
    import traceback

    _SUITE["active"] = True
    all_good = True
    for mod, argv in cases:
        sys.argv = [mod] + [str(a) for a in argv]
        good = False
        try:
            runpy.run_module(mod, run_name="__main__")
        except SystemExit as e:
            good = (e.code in (0, None))
        except Exception:
            traceback.print_exc()
            good = False
        good = all_ok(good) if dist.is_initialized() else good
        dist_print(f"[suite] {mod.split('.')[-1]} {' '.join(str(a) for a in argv)}: {'ok' if good else 'FAILED'}")
        all_good = all_good and good
        if not good:
            break  # peers may be out of step after a failure: stop here
    _SUITE["active"] = False
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier()
    dist_print("PASS" if all_good else "FAIL")
    finalize_distributed()
    return all_good


def bench(fn, args, group=None, flush_l2=True):
    """Device time of ``fn`` in ms (max over ranks). With ``--profile`` the timed loop also runs under torch.profiler and
    the per-rank chrome traces are merged on rank 0 (``prof/<script>/merged.json.gz``), as every reference test does."""
    graph = getattr(args, "graph", False)
    if getattr(args, "profile", False):
        from triton_dist.utils import group_profile

        name = os.path.splitext(os.path.basename(sys.argv[0] or "bench"))[0].split(".")[-1]
        with group_profile(name, do_prof=True, group=group):
            return perf_func_max_over_ranks(fn, args.iters, args.warmup, group=group, flush_l2=flush_l2 and not graph,
                                            use_graph=graph)[1]
    return perf_func_max_over_ranks(fn, args.iters, args.warmup, group=group, flush_l2=flush_l2 and not graph,
                                    use_graph=graph)[1]
