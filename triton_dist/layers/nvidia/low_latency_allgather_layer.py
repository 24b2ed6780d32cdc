"""AllGatherLayer: in-place low-latency all-gather of a symmetric buffer (reference
``layers/nvidia/low_latency_allgather_layer.py:30-200``). All hierarchy variants share the flat
NVSwitch kernels of ``kernels/nvidia/low_latency_allgather.py``."""
from __future__ import annotations

import torch

from ...kernels.nvidia import low_latency_allgather as _ag


class AllGatherLayer:

    def __init__(self, nnodes: int, world_size: int, rank: int, max_buffer_size: int = 2 * 32 * 128 * 128, stages: int = 2):
        self.rank, self.num_ranks, self.nnodes, self.stages = rank, world_size, nnodes, stages
        self.max_buffer_size = max_buffer_size
        self.ctx = _ag.create_fast_allgather_context(rank, 0, world_size, nnodes, max_buffer_size=max(max_buffer_size, 4096))

    @property
    def signal_target(self) -> int:
        return self.ctx.signal_target

    def finalize(self):
        self.ctx.finalize()

    def forward_pull(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_pull(self.ctx, symm_buffer)

    def forward_push_2d(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_2d(self.ctx, symm_buffer)

    def forward_push_3d(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_3d(self.ctx, symm_buffer)

    def forward_push_3d_ll(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_2d_ll(self.ctx, symm_buffer)

    def forward_push_2d_ll(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_2d_ll(self.ctx, symm_buffer)

    def forward_push_2d_ll_multimem(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_2d_ll_multimem(self.ctx, symm_buffer)

    def forward_push_numa_2d(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_numa_2d(self.ctx, symm_buffer)

    def forward_push_numa_2d_ll(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_numa_2d_ll(self.ctx, symm_buffer)

    def forward_push_numa_2d_ll_multinode(self, symm_buffer: torch.Tensor):
        return _ag.fast_allgather_push_numa_2d_ll_multinode(self.ctx, symm_buffer)
