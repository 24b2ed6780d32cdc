"""Megakernel ``add`` task against torch (counterpart of the reference's mega_triton_kernel/test/ops/test_add.py).
Every case builds a one-op graph, runs it twice (monotone scoreboard: the second launch catches stale state) and compares.
    python -m triton_dist.mega_triton_kernel.test.ops.test_add"""
import torch

from triton_dist.mega_triton_kernel import ModelBuilder
from triton_dist.mega_triton_kernel.test.ops.test_ops import check
from triton_dist.test.common import all_ok, finish, setup
from triton_dist.utils import dist_print


def add_case(rank, world, shape, dev, dtype):
    torch.manual_seed(23 + sum(shape))
    a = torch.randn(*shape, device=dev, dtype=dtype)
    b = torch.randn(*shape, device=dev, dtype=dtype)
    out = torch.zeros(*shape, device=dev, dtype=dtype)
    bld = ModelBuilder(rank, world, world)
    bld.make_add(a, b, out)
    bld.compile()
    ok = True
    for it in range(2):
        out.zero_()
        bld.run()
        if dev == "cuda":
            torch.cuda.synchronize()
        ok = check(f"add {tuple(shape)} run{it}", out, (a.float() + b.float()).to(dtype).float(), 1e-6) and ok
        b.add_(1.0)  # the graph holds pointers, not values: new inputs must show up in the next launch
    bld.finalize()
    return ok


if __name__ == "__main__":
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    ok = True
    for shape in [(1, 5120), (8, 5120), (3, 1000), (1, 8), (4, 16384)]:
        ok = add_case(rank, world, shape, dev, dtype) and ok
    ok = all_ok(ok)
    dist_print("[mega add] " + ("PASS" if ok else "FAIL"))
    finish(ok)
