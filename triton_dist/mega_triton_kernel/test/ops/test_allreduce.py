"""Megakernel all-reduce task (NVLS ``multimem.ld_reduce`` of a symmetric partial buffer, residual add folded in) against
``dist.all_reduce`` (counterpart of the reference's mega_triton_kernel/test/ops/test_allreduce.py).
    torchrun --nproc-per-node N -m triton_dist.mega_triton_kernel.test.ops.test_allreduce"""
import torch

from triton_dist.mega_triton_kernel.test.ops.test_ops import allreduce_case
from triton_dist.test.common import all_ok, finish, setup
from triton_dist.utils import dist_print

if __name__ == "__main__":
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    ok = all_ok(allreduce_case(rank, world, dev, dtype))
    dist_print("[mega allreduce] " + ("PASS" if ok else "FAIL"))
    finish(ok)
