"""Small helpers of the megakernel models (reference models/utils.py)."""
import torch


def cos_sin_cache(head_dim: int, max_pos: int, theta: float = 1e6, device=None) -> torch.Tensor:
    """fp32 [max_pos, head_dim] = (cos | sin) table used by the qk-norm + RoPE task."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=device).float() / head_dim))
    ang = torch.arange(max_pos, device=device).float()[:, None] * inv[None, :]
    return torch.cat([ang.cos(), ang.sin()], -1).contiguous()
