"""Eager torch implementations of the megakernel ops (reference mega_triton_kernel/test/torch_impl_utils.py): oracles
for the per-op tests and the bodies of the builder's CPU fallback."""
import torch
import torch.nn.functional as F

from ...kernels.nvidia.flash_decode import torch_paged_decode  # noqa: F401


def torch_rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    xf = x.float()
    return ((xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype).float() * w.float()).to(x.dtype)


def torch_silu_mul_up(x: torch.Tensor) -> torch.Tensor:
    g, u = x.float().chunk(2, -1)
    return (F.silu(g) * u).to(x.dtype)


def torch_linear(x: torch.Tensor, w: torch.Tensor, rms_weight=None, eps: float = 1e-6, fuse_silu: bool = False, bias=None) -> torch.Tensor:
    xin = x if rms_weight is None else torch_rms_norm(x, rms_weight, eps)
    h = F.linear(xin.float(), w.float())
    if fuse_silu:
        n = h.shape[-1] // 2
        h = F.silu(h[..., :n].to(x.dtype).float()) * h[..., n:].to(x.dtype).float()
    if bias is not None:
        h = h + bias.float()
    return h.to(x.dtype)


def torch_rope(x: torch.Tensor, cos_sin: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
    """x [..., D] (neox layout), cos_sin [max_pos, D] = (cos | sin), pos broadcastable to x[..., 0]."""
    D = x.shape[-1]
    cs = cos_sin[pos.long()]
    cos, sin = cs[..., :D // 2], cs[..., D // 2:]
    x1, x2 = x.float()[..., :D // 2], x.float()[..., D // 2:]
    return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], -1).to(x.dtype)
