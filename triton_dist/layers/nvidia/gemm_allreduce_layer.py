"""GemmARLayer: row-parallel linear with fused all-reduce (reference
``layers/nvidia/gemm_allreduce_layer.py:33-116``). One kernel launch per forward, no host sync,
CUDA-graph capturable; the returned tensor is valid until the call after next when
``copy_to_local=False`` on the two-shot path (it aliases the symmetric output buffer)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from ...kernels.nvidia.gemm_allreduce import (create_gemm_ar_context, create_ll_gemm_ar_context, gemm_allreduce_op,
                                              low_latency_gemm_allreduce_op)


class GemmARLayer(torch.nn.Module):

    def __init__(self, tp_group, max_M: int, N: int, K: int, input_dtype: torch.dtype, output_dtype: torch.dtype,
                 local_world_size: int = -1, persistent: bool = True, use_ll_kernel: bool = False, copy_to_local: bool = True,
                 NUM_COMM_SMS: int = 8, user_gemm_config=None):
        super().__init__()
        self.tp_group = tp_group
        # no explicit group = the default (world) group, as everywhere else in the layers
        if tp_group is not None:
            self.rank, self.world_size = tp_group.rank(), tp_group.size()
        elif dist.is_initialized():
            self.rank, self.world_size = dist.get_rank(), dist.get_world_size()
        else:
            self.rank, self.world_size = 0, 1
        self.local_world_size = local_world_size if local_world_size != -1 else self.world_size
        self.max_M, self.N, self.K = max_M, N, K
        self.input_dtype, self.output_dtype = input_dtype, output_dtype
        self.use_ll_kernel = use_ll_kernel
        self.copy_to_local = copy_to_local
        self.NUM_COMM_SMS = NUM_COMM_SMS
        self.persistent = persistent
        self.user_gemm_config = user_gemm_config
        self._init_ctx()

    def _init_ctx(self):
        if self.use_ll_kernel:
            self.ctx = create_ll_gemm_ar_context(self.rank, self.world_size, self.local_world_size, self.max_M, self.N,
                                                 self.output_dtype, NUM_COMM_SMS=self.NUM_COMM_SMS)
        else:
            self.ctx = create_gemm_ar_context(None, self.rank, self.world_size, self.local_world_size, self.max_M, self.N,
                                              self.output_dtype, NUM_COMM_SMS=self.NUM_COMM_SMS)

    def finalize(self):
        if self.ctx is not None:
            self.ctx.finalize()
            self.ctx = None

    def forward(self, input: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        assert input.shape[0] <= self.max_M and weight.shape[0] == self.N
        op = low_latency_gemm_allreduce_op if self.use_ll_kernel else gemm_allreduce_op
        out = op(self.ctx, input, weight, copy_to_local=self.copy_to_local)
        if bias is not None:
            out = out + bias
        return out
