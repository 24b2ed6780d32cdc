"""TP_MoE: tensor-parallel mixture-of-experts FFN (reference ``layers/nvidia/tp_moe.py:48-279``).

Experts are replicated across ranks and every expert's FFN dimension is sharded. ``dist_triton_fwd`` takes
token-sharded activations: router (native GEMM) -> top-k softmax -> AG + grouped GEMM (gate|up) -> SiLU*up ->
grouped GEMM + top-k reduce + reduce-scatter. ``torch_fwd`` is the replicated-activation oracle (per-expert loop +
all-reduce, reference test_ag_moe.py:39-72)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from ...kernels.nvidia.allgather_group_gemm import ag_group_gemm, create_ag_group_gemm_context
from ...kernels.nvidia.allreduce import all_reduce, create_allreduce_ctx
from ...kernels.nvidia.elementwise import linear, silu_mul
from ...kernels.nvidia.moe_reduce_rs import create_moe_rs_context, run_moe_reduce_rs
from ...kernels.nvidia.moe_utils import grouped_gemm, moe_align_sort, moe_gather_rows, reduce_topk
from .tp_mlp import _dev, shard_local


class TP_MoE:

    def __init__(self, rank: int = 0, world_size: int = 8, group=None):
        self.rank, self.world_size, self.group = rank, world_size, group
        self.gate = None  # router [E, H]
        self.w1 = None  # [E, 2*I/W, H]  (gate | up)
        self.w2 = None  # [E, H, I/W]
        self.top_k = 0
        self.norm_topk_prob = True
        self.ag_ctx = self.rs_ctx = self.ar_ctx = None
        self.fwd = self.torch_fwd

    # ------------------------------------------------------------------ parameters
    def _init_parameters(self, mlp, verbose: bool = False):
        """HF Qwen3MoeSparseMoeBlock: mlp.gate (router), mlp.experts[i].{gate_proj, up_proj, down_proj}."""
        W, r = self.world_size, self.rank
        self.gate = mlp.gate.weight.detach().to(_dev())
        holder = mlp if hasattr(mlp, "top_k") else mlp.gate  # transformers >= 5 keeps them on the router module
        self.top_k = holder.top_k
        self.norm_topk_prob = getattr(holder, "norm_topk_prob", True)
        experts = mlp.experts
        if hasattr(experts, "gate_up_proj"):  # fused expert tensors (transformers >= 5): [E, 2I, H] / [E, H, I]
            gu = experts.gate_up_proj.detach()
            inter = gu.shape[1] // 2
            g, u = gu[:, :inter], gu[:, inter:]
            self.w1 = torch.cat((shard_local(g, W, 1, r), shard_local(u, W, 1, r)), dim=1).to(_dev())
            self.w2 = shard_local(experts.down_proj.detach(), W, 2, r).to(_dev())
        else:
            w1, w2 = [], []
            for e in experts:
                g = shard_local(e.gate_proj.weight.detach(), W, 0, r)
                u = shard_local(e.up_proj.weight.detach(), W, 0, r)
                w1.append(torch.cat((g, u), dim=0))
                w2.append(shard_local(e.down_proj.weight.detach(), W, 1, r))
            self.w1, self.w2 = torch.stack(w1).to(_dev()), torch.stack(w2).to(_dev())
        self._post_init()

    def _init_random(self, hidden: int, moe_inter: int, num_experts: int, top_k: int, dtype, gen=None, norm_topk_prob=True,
                     std: float = 0.02):
        i_loc = moe_inter // self.world_size
        dev = _dev()
        router_gen = torch.Generator(device=dev).manual_seed(77)  # the router is replicated: same values on all ranks
        self.gate = (torch.randn(num_experts, hidden, device=dev, generator=router_gen) * std).to(dtype)
        self.w1 = (torch.randn(num_experts, 2 * i_loc, hidden, device=dev, generator=gen) * std).to(dtype)
        self.w2 = (torch.randn(num_experts, hidden, i_loc, device=dev, generator=gen) * std).to(dtype)
        self.top_k, self.norm_topk_prob = top_k, norm_topk_prob
        self._post_init()

    def _post_init(self):
        self.num_experts, _, self.K = self.w1.shape
        self.dtype = self.w1.dtype

    # ------------------------------------------------------------------ contexts
    def _init_ctx(self, max_M, **_ignored):
        max_M = (max_M + self.world_size - 1) // self.world_size * self.world_size
        self.ag_ctx = create_ag_group_gemm_context(max_M, self.K, self.num_experts, self.top_k, self.rank, self.world_size,
                                                   dtype=self.dtype)
        self.rs_ctx = create_moe_rs_context(self.rank, self.world_size, self.world_size, max_M, self.K, self.num_experts,
                                            self.top_k, self.dtype)

    def share_ctx(self, other: "TP_MoE"):
        self.ag_ctx, self.rs_ctx, self.ar_ctx = other.ag_ctx, other.rs_ctx, other.ar_ctx
        if hasattr(other, "ar_method"):
            self.ar_method = other.ar_method

    def _init_AR_ctx(self, max_M, method, dtype=torch.bfloat16):
        self.ar_method = method
        self.ar_ctx = create_allreduce_ctx(workspace_nbytes=max(max_M * self.K * dtype.itemsize, 4096) * 2 * max(self.world_size, 2),
                                           rank=self.rank, world_size=self.world_size, local_world_size=self.world_size)

    def _init_gemm_ar_ctx(self, max_M, dtype=torch.bfloat16):
        self._init_AR_ctx(max_M, None, dtype)

    def finalize(self):
        for c in (self.ag_ctx, self.rs_ctx, self.ar_ctx):
            if c is not None:
                c.finalize()
        self.ag_ctx = self.rs_ctx = self.ar_ctx = None

    # ------------------------------------------------------------------ routing
    def route(self, x2: torch.Tensor):
        logits = linear(x2, self.gate).float() if x2.is_cuda else F.linear(x2.float(), self.gate.float())
        probs = torch.softmax(logits, dim=-1)
        w, ids = torch.topk(probs, self.top_k, dim=-1)
        if self.norm_topk_prob:
            w = w / w.sum(dim=-1, keepdim=True)
        return w, ids

    # ------------------------------------------------------------------ forwards
    @torch.inference_mode()
    def torch_fwd(self, x: torch.Tensor) -> torch.Tensor:
        """Replicated activations; local FFN shard per expert; all-reduce."""
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        w, ids = self.route(x2)
        out = torch.zeros_like(x2, dtype=torch.float32)
        for e in range(self.num_experts):
            tok, k = (ids == e).nonzero(as_tuple=True)
            if tok.numel() == 0:
                continue
            h = F.linear(x2[tok], self.w1[e])
            g, u = h.chunk(2, dim=-1)
            y = F.linear((F.silu(g.float()) * u.float()).to(x.dtype), self.w2[e])
            out.index_add_(0, tok, y.float() * w[tok, k, None])
        out = out.to(x.dtype)
        if self.world_size > 1:
            dist.all_reduce(out, group=self.group)
        return out.view(shape)

    @torch.inference_mode()
    def dist_triton_fwd(self, x: torch.Tensor, **_ignored) -> torch.Tensor:
        """Token-sharded activations [M/W, H] -> [M/W, H]."""
        shape = x.shape
        x2 = x.reshape(-1, shape[-1]).contiguous()
        w_loc, ids_loc = self.route(x2)
        h, sort = ag_group_gemm(x2, self.w1, ids_loc, self.ag_ctx)
        h = silu_mul(h)
        # top-k weights of ALL tokens (tiny): plain all-gather, as the reference does (tp_moe.py:255-256)
        if self.world_size > 1:
            w_all = torch.empty((w_loc.shape[0] * self.world_size, self.top_k), dtype=w_loc.dtype, device=w_loc.device)
            dist.all_gather_into_tensor(w_all, w_loc.contiguous(), group=self.group)
        else:
            w_all = w_loc
        out = run_moe_reduce_rs(h, self.w2, sort, w_all, self.rs_ctx)
        return out.view(*shape[:-1], -1)

    @torch.inference_mode()
    def dist_triton_AR_fwd(self, x: torch.Tensor) -> torch.Tensor:
        """Replicated activations, native grouped GEMMs, native all-reduce."""
        shape = x.shape
        x2 = x.reshape(-1, shape[-1]).contiguous()
        w, ids = self.route(x2)
        sort = moe_align_sort(ids, self.num_experts)
        h = silu_mul(grouped_gemm(moe_gather_rows(x2, sort), self.w1, sort))
        y = grouped_gemm(h, self.w2, sort)
        out = reduce_topk(y, sort, w)
        if self.world_size > 1:
            out = all_reduce(out.contiguous(), ctx=self.ar_ctx)
        return out.view(shape)

    dist_triton_gemm_ar_fwd = dist_triton_AR_fwd
