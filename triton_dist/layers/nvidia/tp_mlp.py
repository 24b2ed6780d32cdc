"""TP_MLP: tensor-parallel gated MLP (reference ``layers/nvidia/tp_mlp.py:52-275``).

gate/up are column-parallel (fused into one [2*I/W, H] weight), down is row-parallel. Forward variants:
  torch_fwd                NCCL all-reduce + torch linears (oracle / baseline)
  dist_triton_fwd          AG+GEMM -> SiLU*up -> GEMM+RS on sequence-sharded activations (fused kernels)
  dist_triton_AR_fwd       native linears + native all-reduce
  dist_triton_gemm_ar_fwd  native gate/up linear + fused GEMM+AllReduce (GemmARLayer)
All non-torch paths run native sm_100a kernels end to end (no cuBLAS, no flashinfer)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from ...kernels.allreduce import AllReduceMethod
from ...kernels.nvidia.allgather_gemm import AllGatherGEMMTensorParallelContext, ag_gemm
from ...kernels.nvidia.allreduce import all_reduce, create_allreduce_ctx
from ...kernels.nvidia.elementwise import linear, silu_mul
from ...kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs
from .gemm_allreduce_layer import GemmARLayer


def shard_local(tensor: torch.Tensor, world_size: int, dim: int, local_rank: int) -> torch.Tensor:
    """This rank's equal slice of ``tensor`` along ``dim``."""
    n = tensor.shape[dim]
    if n % world_size != 0:
        raise ValueError(f"dimension {dim} of size {n} is not divisible by world size {world_size}")
    if not 0 <= local_rank < world_size:
        raise ValueError(f"rank {local_rank} out of range for world size {world_size}")
    return tensor.narrow(dim, local_rank * (n // world_size), n // world_size).contiguous()


def _dev() -> str:
    return "cuda" if torch.cuda.is_available() else "cpu"


class TP_MLP:

    def __init__(self, rank: int = 0, world_size: int = 8, group=None):
        self.rank, self.world_size, self.group = rank, world_size, group
        self.act_fn = F.silu
        self.gate_up_proj: Optional[torch.Tensor] = None
        self.down_proj: Optional[torch.Tensor] = None
        self.ag_ctx = self.rs_ctx = self.ar_ctx = self.gemm_ar_ctx = None
        self.fwd = self.torch_fwd

    # ------------------------------------------------------------------ parameters
    def _init_parameters(self, mlp, verbose: bool = False):
        """Shard an HF-style MLP module (gate_proj / up_proj / down_proj, no bias)."""
        assert getattr(mlp.gate_proj, "bias", None) is None, "bias is not supported"
        self._set_weights(mlp.gate_proj.weight.detach(), mlp.up_proj.weight.detach(), mlp.down_proj.weight.detach())
        if verbose:
            print(f"[RANK {self.rank}] MLP gate_up {tuple(self.gate_up_proj.shape)} down {tuple(self.down_proj.shape)}")

    def _set_weights(self, gate: torch.Tensor, up: torch.Tensor, down: torch.Tensor):
        g = shard_local(gate, self.world_size, 0, self.rank)
        u = shard_local(up, self.world_size, 0, self.rank)
        self.gate_up_proj = torch.cat((g, u), dim=0).to(_dev())  # [2*I/W, H]
        self.down_proj = shard_local(down, self.world_size, 1, self.rank).to(_dev())  # [H, I/W]
        self._post_init()

    def _init_random(self, hidden: int, inter: int, dtype: torch.dtype, gen: Optional[torch.Generator] = None, std: float = 0.02):
        """Random-init shard directly on the device (RANDOM_PARAMS path: no full model on the host)."""
        i_loc = inter // self.world_size
        self.gate_up_proj = (torch.randn(2 * i_loc, hidden, device=_dev(), generator=gen) * std).to(dtype)
        self.down_proj = (torch.randn(hidden, i_loc, device=_dev(), generator=gen) * std).to(dtype)
        self._post_init()

    def _post_init(self):
        self.ag_N_per_rank = self.gate_up_proj.shape[0]
        self.K = self.gate_up_proj.shape[1]
        self.dtype = self.gate_up_proj.dtype

    # ------------------------------------------------------------------ contexts
    def _init_ctx(self, max_M, ag_intranode_stream=None, ag_internode_stream=None, BLOCK_M=128, BLOCK_N=0, BLOCK_K=64, stages=4):
        self.ag_ctx = AllGatherGEMMTensorParallelContext(
            N_per_rank=self.ag_N_per_rank, K=self.K, tensor_dtype=self.dtype, rank=self.rank, num_ranks=self.world_size,
            num_local_ranks=self.world_size, max_M=max_M, BLOCK_M=BLOCK_M, BLOCK_N=0, BLOCK_K=BLOCK_K, stages=stages)
        self.rs_ctx = create_gemm_rs_context(max_M=max_M, N=self.K, rank=self.rank, world_size=self.world_size,
                                             local_world_size=self.world_size, output_dtype=self.dtype)

    def _init_AR_ctx(self, max_M, method: AllReduceMethod, dtype=torch.bfloat16):
        self.ar_method = method
        N = self.down_proj.shape[0]
        self.ar_ctx = create_allreduce_ctx(workspace_nbytes=max(max_M * N * dtype.itemsize, 4096) * 2 * max(self.world_size, 2),
                                           rank=self.rank, world_size=self.world_size, local_world_size=self.world_size)

    def _init_gemm_ar_ctx(self, max_M, dtype=torch.bfloat16):
        N, K = self.down_proj.shape
        self.gemm_ar_ctx = GemmARLayer(self.group, max_M, N, K, dtype, dtype, self.world_size, persistent=True,
                                       use_ll_kernel=max_M <= 256, copy_to_local=False)

    def finalize(self):
        for c in (self.ag_ctx, self.rs_ctx, self.ar_ctx, self.gemm_ar_ctx):
            if c is not None:
                c.finalize()
        self.ag_ctx = self.rs_ctx = self.ar_ctx = self.gemm_ar_ctx = None

    # ------------------------------------------------------------------ forwards
    @torch.inference_mode()
    def torch_fwd(self, x: torch.Tensor) -> torch.Tensor:
        out = F.linear(x, self.gate_up_proj)
        wg, w1 = torch.chunk(out, 2, dim=-1)
        out = F.linear(self.act_fn(wg) * w1, self.down_proj)
        if self.world_size > 1:
            dist.all_reduce(out, dist.ReduceOp.SUM, group=self.group)
        return out

    @torch.inference_mode()
    def dist_triton_fwd(self, x: torch.Tensor, ag_gemm_persistent=True, gemm_rs_persistent=True, autotune=False) -> torch.Tensor:
        """x is sharded on tokens: [M/W, H] (or [B/W, S, H]); so is the output."""
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        h = ag_gemm(x2, self.gate_up_proj, ctx=self.ag_ctx)
        h = silu_mul(h)
        out = gemm_rs(h, self.down_proj, self.rs_ctx)
        return out.view(*shape[:-1], -1)

    @torch.inference_mode()
    def dist_triton_AR_fwd(self, x: torch.Tensor) -> torch.Tensor:
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        h = silu_mul(linear(x2, self.gate_up_proj))
        out = linear(h, self.down_proj)
        if self.world_size > 1:
            out = all_reduce(out.contiguous(), method=self.ar_method, ctx=self.ar_ctx)
        return out.view(shape)

    @torch.inference_mode()
    def dist_triton_gemm_ar_fwd(self, x: torch.Tensor) -> torch.Tensor:
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        h = silu_mul(linear(x2, self.gate_up_proj))
        out = self.gemm_ar_ctx.forward(h, self.down_proj)
        return out.view(shape)

    # ------------------------------------------------------------------ op-level helpers (reference :230-275)
    @torch.inference_mode()
    def torch_ag_gemm(self, x: torch.Tensor) -> torch.Tensor:
        from ...kernels.nvidia.allgather_gemm import torch_ag_gemm

        return torch_ag_gemm(x, self.gate_up_proj, self.group)

    @torch.inference_mode()
    def dist_triton_ag_gemm(self, x: torch.Tensor, persistent=True, autotune=False) -> torch.Tensor:
        return ag_gemm(x, self.gate_up_proj, ctx=self.ag_ctx)

    @torch.inference_mode()
    def torch_gemm_rs(self, x: torch.Tensor) -> torch.Tensor:
        from ...kernels.nvidia.gemm_reduce_scatter import torch_gemm_rs

        return torch_gemm_rs(x, self.down_proj, self.group)

    @torch.inference_mode()
    def dist_triton_gemm_rs(self, x: torch.Tensor, persistent=True) -> torch.Tensor:
        return gemm_rs(x, self.down_proj, self.rs_ctx)
