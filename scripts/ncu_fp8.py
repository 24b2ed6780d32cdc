"""Tiny driver for the ncu capture of the fp8 block-scaled GEMM (one shape, a few launches)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.kernels.nvidia.gemm_fp8 import gemm_fp8, quant_fp8_1x128, quant_fp8_128x128

a8, sa = quant_fp8_1x128(torch.randn(8192, 7168, device="cuda"))
b8, sb = quant_fp8_128x128(torch.randn(4096, 7168, device="cuda") * 0.05)
out = torch.empty(8192, 4096, device="cuda", dtype=torch.bfloat16)
for _ in range(5):
    gemm_fp8(a8, sa, b8, sb, out=out)
torch.cuda.synchronize()
print("ok")
