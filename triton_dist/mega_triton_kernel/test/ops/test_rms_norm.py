"""Megakernel ``rms_norm`` task (plain, and with the residual add folded in) against an fp32 torch formula (counterpart
of the reference's mega_triton_kernel/test/ops/test_rms_norm.py).
    python -m triton_dist.mega_triton_kernel.test.ops.test_rms_norm"""
import torch

from triton_dist.mega_triton_kernel import ModelBuilder
from triton_dist.mega_triton_kernel.test.ops.test_ops import check
from triton_dist.test.common import all_ok, finish, setup
from triton_dist.utils import dist_print


def rms_case(rank, world, M, H, with_residual, dev, dtype, eps=1e-6):
    torch.manual_seed(31 + M + H)
    x = torch.randn(M, H, device=dev, dtype=dtype)
    w = (torch.rand(H, device=dev) + 0.5).to(dtype)
    res0 = torch.randn(M, H, device=dev, dtype=dtype) if with_residual else None
    res = res0.clone() if with_residual else None
    y = torch.zeros(M, H, device=dev, dtype=dtype)
    bld = ModelBuilder(rank, world, world)
    bld.make_rms_norm(x, w, y, eps, residual=res)
    bld.compile()
    ok = True
    acc = res0.float() if with_residual else None
    for it in range(2):
        bld.run()
        if dev == "cuda":
            torch.cuda.synchronize()
        if with_residual:  # residual <- x + residual (rounded to the storage type), output = norm of the new residual
            acc = (x.float() + acc).to(dtype).float()
            xs = acc
            ok = check(f"rms_norm residual M={M} H={H} run{it}", res, acc, 1e-6) and ok
        else:
            xs = x.float()
        ref = (xs * torch.rsqrt(xs.pow(2).mean(-1, keepdim=True) + eps) * w.float())
        ok = check(f"rms_norm M={M} H={H} residual={with_residual} run{it}", y, ref, 2e-2) and ok
    bld.finalize()
    return ok


if __name__ == "__main__":
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    ok = True
    for M, H, r in [(1, 5120, False), (1, 5120, True), (8, 4096, True), (3, 1024, False), (2, 8192, True)]:
        ok = rms_case(rank, world, M, H, r, dev, dtype) and ok
    ok = all_ok(ok)
    dist_print("[mega rms_norm] " + ("PASS" if ok else "FAIL"))
    finish(ok)
