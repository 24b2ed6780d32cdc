"""Tiny driver for ncu captures: a few launches of the plain GEMM (single-CTA and CTA-pair kernels) at the flagship
per-GPU shape and one split-KV flash decode + megakernel-style GEMV."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.kernels.nvidia import gemm_persistent
from triton_dist.kernels.nvidia.elementwise import gemv

M, N, K = 8192, 4096, 4096
a = (torch.rand(M, K, device="cuda") * 2 - 1).bfloat16()
b = ((torch.rand(N, K, device="cuda") * 2 - 1) * 0.1).bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    gemm_persistent(a, b, out=out, block_n=256, cta_group=1)
for _ in range(3):
    gemm_persistent(a, b, out=out, block_n=256, cta_group=2)
x = torch.randn(1, 5120, device="cuda").bfloat16()
w = (0.05 * torch.randn(6400, 5120, device="cuda")).bfloat16()
for _ in range(3):
    gemv(x, w)
torch.cuda.synchronize()
print("ok")
