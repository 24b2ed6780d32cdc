"""Tiny driver for the ncu capture of the tcgen05 flash-attention forward."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.kernels.nvidia.sp_ag_attention_intra_node import flash_attn_varlen

B, L, Hq, Hkv = 2, 4096, 32, 8
q = torch.randn(B * L, Hq, 128, device="cuda").bfloat16()
k = torch.randn(B * L, Hkv, 128, device="cuda").bfloat16()
v = torch.randn(B * L, Hkv, 128, device="cuda").bfloat16()
cu = torch.arange(B + 1, device="cuda", dtype=torch.int32) * L
for _ in range(4):
    flash_attn_varlen(q, k, v, cu, cu, L, causal=False)
torch.cuda.synchronize()
print("ok")
