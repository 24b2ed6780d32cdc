"""EPAll2AllLayer: expert-parallel token dispatch / combine (reference ``layers/nvidia/ep_a2a_layer.py:40-248``).

``dispatch(input, exp_indices)`` delivers every (token, k) pair to the rank that owns expert ``exp_indices[t, k]`` and
returns the received rows grouped by local expert; ``combine(expert_out)`` brings the expert outputs back and sums
the top-k contributions of every token (optionally weighted). Unlike the reference there is no pinned-memory host
polling: buffers have a fixed capacity, counts stay on the device, and both calls are CUDA-graph capturable."""
from __future__ import annotations

from typing import Optional

import torch

from ...kernels.nvidia.ep_a2a import EPContext, ep_combine, ep_dispatch


class EPAll2AllLayer(torch.nn.Module):

    def __init__(self, ep_group, max_tokens: int, hidden: int, topk: int, rank: int, num_tot_experts: int, local_world_size: int,
                 world_size: int, dtype=torch.bfloat16, num_sm: int = 32, fp8_dispatch: bool = False, capacity: int = 0):
        super().__init__()
        assert num_tot_experts % world_size == 0
        assert local_world_size == world_size, "single NVSwitch node only"
        self.ep_group, self.rank, self.world_size = ep_group, rank, world_size
        self.max_tokens, self.hidden, self.topk, self.dtype = max_tokens, hidden, topk, dtype
        self.num_tot_experts = num_tot_experts
        self.experts_per_rank = num_tot_experts // world_size
        self.ctx = EPContext(rank=rank, world_size=world_size, max_tokens=max_tokens, hidden=hidden, topk=topk,
                             num_experts=num_tot_experts, dtype=dtype, fp8=fp8_dispatch, capacity=capacity, num_sm=num_sm)
        self.num_dispatch_token_cur_rank = None

    def finalize(self):
        self.ctx.finalize()

    def dispatch(self, input: torch.Tensor, exp_indices: torch.Tensor, exact: bool = True):
        """Returns the received rows. ``exact=True`` (reference behaviour) trims to the received count, which costs one
        device->host read; ``exact=False`` returns ``(buffer, expert_counts, num_recv)`` with everything on the device."""
        self.num_dispatch_token_cur_rank = input.shape[0]
        buf, counts, n = ep_dispatch(self.ctx, input, exp_indices)
        if not exact:
            return buf, counts, n
        return buf[:int(n.item())].clone()

    def dispatch_scales(self) -> torch.Tensor:
        """fp8 dispatch: per-row fp32 scales [capacity, hidden / 128] matching the last dispatch."""
        return self.ctx.recv_scale

    def recv_expert_counts(self) -> torch.Tensor:
        return self.ctx.recv_expert_counts

    def combine(self, input: torch.Tensor, topk_weights: Optional[torch.Tensor] = None) -> torch.Tensor:
        return ep_combine(self.ctx, input, topk_weights)
