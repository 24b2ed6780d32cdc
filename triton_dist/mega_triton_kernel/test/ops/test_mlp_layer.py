"""Megakernel linear / MLP tasks against torch (counterpart of the reference's mega_triton_kernel/test/ops/test_mlp_layer.py):
the single-linear shapes (norm prologue, SiLU pairing, bias, N not divisible by the CTA count, K up to 25600) and the chained
embedding -> fc1 -> fc2 -> add -> LM head -> argmax graph of test_ops.py.
    python -m triton_dist.mega_triton_kernel.test.ops.test_mlp_layer"""
import torch

from triton_dist.mega_triton_kernel.test.ops.test_ops import chain_case, linear_case
from triton_dist.test.common import all_ok, finish, setup
from triton_dist.utils import dist_print

if __name__ == "__main__":
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    ok = True
    for case in [(1, 3200, 5120, True, True, False), (1, 5120, 3200, False, False, False), (4, 1280, 5120, True, False, True),
                 (8, 296, 2048, True, True, True)]:
        ok = linear_case(rank, world, *case, dev, dtype) and ok
    ok = chain_case(rank, world, dev, dtype) and ok
    ok = all_ok(ok)
    dist_print("[mega mlp_layer] " + ("PASS" if ok else "FAIL"))
    finish(ok)
