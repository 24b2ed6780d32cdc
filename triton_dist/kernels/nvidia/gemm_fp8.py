"""Block-scaled fp8 (e4m3) GEMM, plain and grouped (MoE), on ``tcgen05.mma.kind::f8f6f4``.

No counterpart in the reference: its grouped GEMMs are bf16/fp16 (``kernels/nvidia/moe_reduce_rs.py:69``) and fp8 exists
only as the all-to-all payload plus the tutorial quantiser (``tutorials/04-deepseek-infer-all2all.py:334-427``). Here the
fp8 rows and 1 x 128 scales that the EP dispatch delivers are multiplied as they are (csrc/kernels/gemm_fp8_sm100.cuh):
activations carry one fp32 scale per row and 128-wide K block, weights one per 128 x 128 block (the DeepSeek-V3 recipe).
Every function has a plain-torch path for CPU runs, which is also the oracle of the tests.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ... import _lib
from ...utils import gemm_metadata, launch_metadata

FP8 = torch.float8_e4m3fn
FP8_MAX = 448.0
BLOCK = 128


def quant_fp8_1x128(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """[M, K] -> (e4m3 [M, K], fp32 scales [M, ceil(K/128)]): x ~ x8 * scale, one scale per row and 128 columns."""
    M, K = x.shape
    kb = (K + BLOCK - 1) // BLOCK
    xp = torch.nn.functional.pad(x.float(), (0, kb * BLOCK - K)).view(M, kb, BLOCK)
    scale = (xp.abs().amax(-1) / FP8_MAX).clamp_min(1e-12)
    x8 = (xp / scale[..., None]).clamp(-FP8_MAX, FP8_MAX).view(M, kb * BLOCK)[:, :K].to(FP8)
    return x8.contiguous(), scale.contiguous()


def quant_fp8_128x128(w: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """[..., N, K] -> (e4m3 [..., N, K], fp32 scales [..., ceil(N/128), ceil(K/128)]): one scale per 128 x 128 block."""
    *lead, N, K = w.shape
    nb, kb = (N + BLOCK - 1) // BLOCK, (K + BLOCK - 1) // BLOCK
    wp = torch.nn.functional.pad(w.float(), (0, kb * BLOCK - K, 0, nb * BLOCK - N)).view(*lead, nb, BLOCK, kb, BLOCK)
    scale = (wp.abs().amax(dim=(-3, -1)) / FP8_MAX).clamp_min(1e-12)
    w8 = (wp / scale[..., :, None, :, None]).clamp(-FP8_MAX, FP8_MAX).view(*lead, nb * BLOCK, kb * BLOCK)[..., :N, :K].to(FP8)
    return w8.contiguous(), scale.contiguous()


def dequant_1x128(x8: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    K = x8.shape[-1]
    return x8.float() * scale.repeat_interleave(BLOCK, dim=-1)[..., :K]


def dequant_128x128(w8: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    N, K = w8.shape[-2:]
    s = scale.repeat_interleave(BLOCK, dim=-2).repeat_interleave(BLOCK, dim=-1)[..., :N, :K]
    return w8.float() * s


@launch_metadata("gemm_fp8", lambda a8, sa, b8, *_, **__: gemm_metadata(a8.shape[0], b8.shape[-2], a8.shape[1], 1))
def gemm_fp8(a8: torch.Tensor, sa: torch.Tensor, b8: torch.Tensor, sb: torch.Tensor, out: Optional[torch.Tensor] = None,
             out_dtype: torch.dtype = torch.bfloat16, block_expert: Optional[torch.Tensor] = None, num_sms: int = 0) -> torch.Tensor:
    """out[M, N] = dequant(a8, sa) @ dequant(b8, sb)^T.

    a8 [M, K] e4m3, sa [M, ceil(K/128)] fp32; b8 [N, K] (plain) or [E, N, K] (grouped), sb [ceil(N/128), ceil(K/128)] or
    [E, ...]. Grouped: ``block_expert`` [M/128] int32 names the expert of every 128-row block of a8 (negative = padding
    block, left untouched)."""
    M, K = a8.shape
    grouped = block_expert is not None
    E = b8.shape[0] if grouped else 1
    N = b8.shape[-2]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a8.device)
    if not (a8.is_cuda and _lib.use_native()):
        a = dequant_1x128(a8, sa)
        if not grouped:
            out.copy_((a @ dequant_128x128(b8, sb).t()).to(out.dtype))
        else:
            out.zero_()
            w = dequant_128x128(b8, sb)
            for blk in range(M // BLOCK):
                e = int(block_expert[blk])
                if e >= 0:
                    out[blk * BLOCK:(blk + 1) * BLOCK] = (a[blk * BLOCK:(blk + 1) * BLOCK] @ w[e].t()).to(out.dtype)
        return out
    assert a8.dtype == FP8 and b8.dtype == FP8 and a8.stride(-1) == 1 and b8.is_contiguous()
    assert sa.dtype == torch.float32 and sb.dtype == torch.float32 and sa.stride(-1) == 1 and sb.is_contiguous()
    assert sa.shape[0] == M and sa.shape[1] >= (K + BLOCK - 1) // BLOCK
    _lib.check(_lib.lib().tdist_gemm_fp8(a8.data_ptr(), sa.data_ptr(), sa.stride(0), b8.data_ptr(), sb.data_ptr(), out.data_ptr(),
                                         M, N, K, a8.stride(0), K, out.stride(0), _lib.dtype_code(out.dtype),
                                         block_expert.data_ptr() if grouped else None, E, num_sms, _lib.stream_ptr()),
               "gemm_fp8")
    return out
