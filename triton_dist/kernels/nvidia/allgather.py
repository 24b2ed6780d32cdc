"""All-gather producers. API of ``kernels/nvidia/allgather.py`` in the reference
(method enum :46-71, copy-engine producers :81-290, inter-node :293-560).

On one NVSwitch node every topology-specific variant (ring 1-D / NUMA 2-D /
inter-node) maps to the same one-hop exchange, so the enum is kept for API
compatibility and resolves to ``All2All_IntraNode``. The copy-engine producer
(``cudaMemcpyAsync`` per shard + stream write of a flag) is kept as the
baseline path; the fused ops do not use it — their gather runs inside the GEMM
kernel.
"""
from __future__ import annotations

import functools
from enum import Enum
from typing import List, Optional

import torch
import torch.distributed as dist

from ... import _lib
from ...runtime import symm as _symm
from ..common_ops import set_signal


class AllGatherMethod(Enum):
    Auto = 0
    All2All_IntraNode = 1
    All2All_InterNode = 2
    Ring1D_IntraNode = 3
    Ring2D_IntraNode = 4
    Ring1D_InterNode = 5
    Ring2D_InterNode = 6


@functools.lru_cache()
def get_auto_all_gather_method(num_ranks: int, num_local_ranks: int) -> AllGatherMethod:
    if num_ranks != num_local_ranks:
        return AllGatherMethod.All2All_InterNode
    return AllGatherMethod.All2All_IntraNode  # NVSwitch: flat full mesh


def cp_engine_producer_all_gather_intra_node(rank: int, num_ranks: int, local_tensor: torch.Tensor,
                                             remote_tensor_buffers: List[torch.Tensor], ag_stream: "torch.cuda.Stream",
                                             signal_buffers: Optional[List[torch.Tensor]] = None, signal_target: int = 1,
                                             all_gather_method: AllGatherMethod = AllGatherMethod.Auto,
                                             for_correctness: bool = False) -> None:
    """Full-mesh *pull* by the copy engines on ``ag_stream``: for every source rank copy its shard
    from that rank's symmetric buffer into mine, then publish ``signal[src] = signal_target``.

    ``remote_tensor_buffers[p]`` is peer p's [num_ranks * M_per_rank, K] buffer (peer views of one
    symmetric tensor), each rank having already placed its own shard in slot ``rank``.
    """
    m = local_tensor.shape[0]
    mine = remote_tensor_buffers[rank]
    with torch.cuda.stream(ag_stream):
        for i in range(1, num_ranks):
            src = (rank + i) % num_ranks
            if for_correctness:
                torch.cuda._sleep(int(torch.randint(0, 1_000_000, (1,)).item()))
            mine[src * m:(src + 1) * m].copy_(remote_tensor_buffers[src][src * m:(src + 1) * m], non_blocking=True)
            if signal_buffers is not None:
                set_signal(signal_buffers[rank][src:src + 1], signal_target, ag_stream, use_memops=True)


def cp_engine_producer_all_gather_inter_node(*args, **kwargs):
    raise NotImplementedError("single NVSwitch node only: there is no inter-node path on this platform")


def torch_all_gather(x: torch.Tensor, group=None) -> torch.Tensor:
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return x.clone()
    out = torch.empty((world * x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x.contiguous(), group=group)
    return out
