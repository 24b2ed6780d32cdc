"""Tiny driver for an ncu capture of the megakernel LINEAR task: one op (gate_up-shaped, silu pairing), 3 launches."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.mega_triton_kernel import ModelBuilder

N, K = 25600, 5120
w = torch.randn(2 * N, K, device="cuda", dtype=torch.bfloat16) * 0.02
x = torch.randn(1, K, device="cuda", dtype=torch.bfloat16)
y = torch.zeros(1, N, device="cuda", dtype=torch.bfloat16)
b = ModelBuilder(0, 1, 1)
b.make_linear(x, w, y, fuse_silu=True)
b.compile()
for _ in range(3):
    b.run()
torch.cuda.synchronize()
print("ok")
