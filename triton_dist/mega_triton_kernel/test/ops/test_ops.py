"""Per-op checks of the megakernel task bodies against plain PyTorch (fp32 reference of the same op).

Counterpart of the reference's mega_triton_kernel/test/ops/{test_mlp_layer,test_rms_norm,test_add,test_allreduce,
test_page_attn}.py, folded into one script: every case builds a tiny graph with ModelBuilder, runs it TWICE (the
scoreboard and the weight ring are monotone across launches, the second run catches stale state) and compares.
    torchrun --nproc-per-node N -m triton_dist.mega_triton_kernel.test.ops.test_ops
"""
import argparse

import torch
import torch.distributed as dist
import torch.nn.functional as F

from triton_dist.mega_triton_kernel import ModelBuilder
from triton_dist.test.common import all_ok, finish, setup
from triton_dist.utils import dist_print, nvshmem_create_tensor, nvshmem_free_tensor_sync


def _rms(x, w, eps):
    xf = x.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype).float() * w.float()


def check(name, got, ref, tol):
    err = (got.float() - ref).abs().max().item()
    scale = max(ref.abs().max().item(), 1e-6)
    good = err <= tol * scale
    if not good:
        dist_print(f"[mega ops] {name}: max abs err {err:.4g} vs scale {scale:.4g} FAILED", allowed_ranks="all", prefix=True)
    return good


def linear_case(rank, world, M, N, K, norm, pair, bias, dev, dtype):
    torch.manual_seed(17 + M + N + K)
    x = torch.randn(M, K, device=dev, dtype=dtype)
    w = torch.randn((2 * N if pair else N), K, device=dev, dtype=dtype) * (K ** -0.5)
    nw = (torch.rand(K, device=dev, dtype=dtype) + 0.5) if norm else None
    b = torch.randn(N, device=dev, dtype=dtype) if bias else None
    y = torch.zeros(M, N, device=dev, dtype=dtype)
    bld = ModelBuilder(rank, world, world)
    bld.make_linear(x, w, y, rms_weight=nw, rms_eps=1e-6, fuse_silu=pair, bias=b)
    bld.compile()
    xin = _rms(x, nw, 1e-6).to(dtype).float() if norm else x.float()
    h = xin @ w.float().t()
    if pair:
        g, u = h[:, :N].to(dtype).float(), h[:, N:].to(dtype).float()
        h = F.silu(g) * u
    if bias:
        h = h + b.float()
    ok = True
    for it in range(2):
        y.zero_()
        bld.run()
        torch.cuda.synchronize() if dev == "cuda" else None
        ok = check(f"linear M={M} N={N} K={K} norm={norm} pair={pair} bias={bias} run{it}", y, h, 2e-2) and ok
    bld.finalize()
    return ok


def chain_case(rank, world, dev, dtype):
    """norm -> fc1(silu pair) -> fc2 -> add, plus embedding/argmax: exercises dependencies between linear tasks and the
    weight stream running ahead across tasks."""
    torch.manual_seed(5)
    M, H, I, V = 2, 1024, 2816, 1000
    table = torch.randn(V, H, device=dev, dtype=dtype)
    ids = torch.tensor([3, 77], device=dev, dtype=torch.long)
    x = torch.zeros(M, H, device=dev, dtype=dtype)
    nw = torch.rand(H, device=dev, dtype=dtype) + 0.5
    w1 = torch.randn(2 * I, H, device=dev, dtype=dtype) * (H ** -0.5)
    w2 = torch.randn(H, I, device=dev, dtype=dtype) * (I ** -0.5)
    wl = torch.randn(V, H, device=dev, dtype=dtype) * (H ** -0.5)
    act = torch.zeros(M, I, device=dev, dtype=dtype)
    h2 = torch.zeros(M, H, device=dev, dtype=dtype)
    res = torch.zeros(M, H, device=dev, dtype=dtype)
    logits = torch.zeros(M, V, device=dev, dtype=dtype)
    nxt = torch.zeros(M, device=dev, dtype=torch.long)
    bld = ModelBuilder(rank, world, world)
    bld.make_embedding(ids, table, x)
    bld.make_fc1(x, w1, act, rms_weight=nw, fuse_silu=True)
    bld.make_fc2(act, w2, h2)
    bld.make_add(h2, x, res)
    bld.make_linear(res, wl, logits, rms_weight=nw)
    bld.make_argmax(logits, nxt)
    bld.compile()
    xr = table[ids].float()
    hh = _rms(xr.to(dtype), nw, 1e-6).to(dtype).float() @ w1.float().t()
    a = (F.silu(hh[:, :I].to(dtype).float()) * hh[:, I:].to(dtype).float()).to(dtype).float()
    r = ((a @ w2.float().t()).to(dtype).float() + xr).to(dtype)
    lg = _rms(r, nw, 1e-6).to(dtype).float() @ wl.float().t()
    ok = True
    for it in range(3):
        bld.run()
        torch.cuda.synchronize() if dev == "cuda" else None
        ok = check(f"chain logits run{it}", logits, lg, 3e-2) and ok
        ok = bool(torch.equal(nxt, logits.float().argmax(-1))) and ok
    bld.finalize()
    return ok


def allreduce_case(rank, world, dev, dtype):
    torch.manual_seed(11 + rank)
    n = 2 * 5120
    src = nvshmem_create_tensor((n,), dtype)
    src.copy_(torch.randn(n, device=dev, dtype=dtype))
    res = torch.ones(n, device=dev, dtype=dtype)
    out = torch.zeros(n, device=dev, dtype=dtype)
    bld = ModelBuilder(rank, world, world)
    bld.make_allreduce(src, out, residual=res)
    bld.compile()
    tot = src.float().clone()
    if world > 1:
        dist.all_reduce(tot)
    ok = True
    for it in range(2):
        bld.run()
        torch.cuda.synchronize() if dev == "cuda" else None
        ok = check(f"allreduce+residual run{it}", out, tot * (it + 1) + 1.0, 2e-2) and ok
    bld.finalize()
    nvshmem_free_tensor_sync(src)
    return ok


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    args = ap.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    ok = True
    # (M, N, K, norm, pair, bias): K/8 multiple of 32 or not (row-straddling vector groups), tiles smaller and larger
    # than one 16 KB chunk, N not divisible by the CTA count, every batch template
    for case in [(1, 640, 1024, False, False, False), (1, 5120, 3200, True, False, False), (2, 1280, 5120, True, False, True),
                 (1, 3200, 5120, True, True, False), (4, 1000, 320, False, False, False), (8, 296, 2048, True, True, True),
                 (3, 18992, 1024, True, False, False), (1, 5120, 25600, False, False, False), (1, 37, 8192, False, False, False)]:
        ok = linear_case(rank, world, *case, dev, dtype) and ok
    ok = chain_case(rank, world, dev, dtype) and ok
    ok = allreduce_case(rank, world, dev, dtype) and ok
    ok = all_ok(ok)
    dist_print("[mega ops] " + ("PASS" if ok else "FAIL"))
    finish(ok)
