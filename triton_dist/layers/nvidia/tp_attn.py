"""TP_Attn: tensor-parallel GQA attention block (reference ``layers/nvidia/tp_attn.py:79-324``).

wq/wk/wv are column-parallel over heads (fused into wqkv), wo is row-parallel. After the QKV projection one
native kernel applies the per-head q/k RMSNorm (Qwen3), RoPE and writes K/V straight into the KV cache; decode
steps (q_len == 1) then run the native split-KV flash-decode over the cache. Prefill attention uses torch SDPA
(the reference uses the flash-attn library at the same place).

Forward variants mirror the reference: torch_fwd | dist_triton_fwd (AG+GEMM / GEMM+RS, token-sharded
activations) | dist_triton_AR_fwd | dist_triton_gemm_ar_fwd."""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from ...kernels.allreduce import AllReduceMethod
from ...kernels.nvidia.allgather_gemm import AllGatherGEMMTensorParallelContext, ag_gemm
from ...kernels.nvidia.allreduce import all_reduce, create_allreduce_ctx
from ...kernels.nvidia.elementwise import linear, make_cos_sin_cache, qk_norm_rope_update_kvcache, rmsnorm
from ...kernels.nvidia.flash_decode import dense_decode
from ...kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs
from .gemm_allreduce_layer import GemmARLayer
from .tp_mlp import _dev, shard_local


def layer_norm(hidden_states: torch.Tensor, eps: float, w: torch.Tensor) -> torch.Tensor:
    """RMSNorm over the last dim (name kept from the reference, tp_attn.py:60-66)."""
    return rmsnorm(hidden_states, w, eps)


def _set_cos_sin_cache(inv_freq: torch.Tensor, max_length: int) -> torch.Tensor:
    """fp32 [max_length, 128] = cos(64) | sin(64) (reference tp_attn.py:69-76)."""
    t = torch.arange(max_length, device=inv_freq.device, dtype=torch.float32)
    freqs = torch.outer(t, inv_freq.float())
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).contiguous()


class TP_Attn:

    def __init__(self, rank: int = 0, world_size: int = 8, group=None, head_dim: int = 128):
        self.rank, self.world_size, self.group = rank, world_size, group
        self.head_dim = head_dim
        self.wqkv = self.wo = None
        self.q_norm_w = self.k_norm_w = None
        self.q_norm_eps = self.k_norm_eps = 1e-6
        self.bqkv = None
        self.ag_ctx = self.rs_ctx = self.ar_ctx = self.gemm_ar_ctx = None
        self.fwd = self.torch_fwd

    # ------------------------------------------------------------------ parameters
    def _init_parameters(self, self_attn, verbose: bool = False):
        W, r, D = self.world_size, self.rank, self.head_dim
        wq, wk, wv = (m.weight.detach() for m in (self_attn.q_proj, self_attn.k_proj, self_attn.v_proj))
        bias = None
        if getattr(self_attn.q_proj, "bias", None) is not None:
            bias = tuple(m.bias.detach() for m in (self_attn.q_proj, self_attn.k_proj, self_attn.v_proj))
        qn = getattr(self_attn, "q_norm", None)
        kn = getattr(self_attn, "k_norm", None)
        self._set_weights(wq, wk, wv, self_attn.o_proj.weight.detach(), bias,
                          None if qn is None else (qn.weight.detach(), qn.variance_epsilon),
                          None if kn is None else (kn.weight.detach(), kn.variance_epsilon))
        if verbose:
            print(f"[RANK {self.rank}] Attn qkv {tuple(self.wqkv.shape)} o {tuple(self.wo.shape)}")

    def _kv_shard(self, t: torch.Tensor) -> torch.Tensor:
        """KV heads are sharded; when there are fewer KV heads than ranks each head is replicated."""
        W, r, D = self.world_size, self.rank, self.head_dim
        n_kv = t.shape[0] // D
        if n_kv >= W:
            return shard_local(t, W, 0, r)
        head = r * n_kv // W
        return t.narrow(0, head * D, D).contiguous()

    def _set_weights(self, wq, wk, wv, wo, bias=None, q_norm=None, k_norm=None):
        W, r = self.world_size, self.rank
        q = shard_local(wq, W, 0, r)
        k, v = self._kv_shard(wk), self._kv_shard(wv)
        self.q_size, self.kv_size = q.shape[0], k.shape[0]
        self.wqkv = torch.cat((q, k, v), dim=0).to(_dev())
        self.wo = shard_local(wo, W, 1, r).to(_dev())
        if bias is not None:
            self.bqkv = torch.cat((shard_local(bias[0], W, 0, r), self._kv_shard(bias[1]), self._kv_shard(bias[2]))).to(_dev())
        if q_norm is not None:
            self.q_norm_w, self.q_norm_eps = q_norm[0].to(_dev()), q_norm[1]
        if k_norm is not None:
            self.k_norm_w, self.k_norm_eps = k_norm[0].to(_dev()), k_norm[1]
        self._post_init()

    def _init_random(self, hidden: int, num_heads: int, num_kv_heads: int, dtype, qk_norm: bool = True, qkv_bias: bool = False,
                     gen: Optional[torch.Generator] = None, std: float = 0.02):
        W, D = self.world_size, self.head_dim
        self.q_size = num_heads // W * D
        self.kv_size = max(num_kv_heads // W, 1) * D
        dev = _dev()
        self.wqkv = (torch.randn(self.q_size + 2 * self.kv_size, hidden, device=dev, generator=gen) * std).to(dtype)
        self.wo = (torch.randn(hidden, self.q_size, device=dev, generator=gen) * std).to(dtype)
        if qk_norm:
            self.q_norm_w = torch.ones(D, device=dev, dtype=dtype)
            self.k_norm_w = torch.ones(D, device=dev, dtype=dtype)
        if qkv_bias:
            self.bqkv = torch.zeros(self.q_size + 2 * self.kv_size, device=dev, dtype=dtype)
        self._post_init()

    def _post_init(self):
        self.ag_N_per_rank, self.K = self.wqkv.shape
        self.dtype = self.wqkv.dtype
        self.num_q_heads = self.q_size // self.head_dim
        self.num_kv_heads = self.kv_size // self.head_dim

    # ------------------------------------------------------------------ contexts
    def _init_ctx(self, max_M, ag_intranode_stream=None, ag_internode_stream=None, BLOCK_M=128, BLOCK_N=0, BLOCK_K=64, stages=4):
        self.ag_ctx = AllGatherGEMMTensorParallelContext(
            N_per_rank=self.ag_N_per_rank, K=self.K, tensor_dtype=self.dtype, rank=self.rank, num_ranks=self.world_size,
            num_local_ranks=self.world_size, max_M=max_M, BLOCK_M=BLOCK_M, BLOCK_N=0, BLOCK_K=BLOCK_K, stages=stages)
        self.rs_ctx = create_gemm_rs_context(max_M=max_M, N=self.K, rank=self.rank, world_size=self.world_size,
                                             local_world_size=self.world_size, output_dtype=self.dtype)

    def _init_AR_ctx(self, max_M, method: AllReduceMethod, dtype=torch.bfloat16):
        self.ar_method = method
        self.ar_ctx = create_allreduce_ctx(workspace_nbytes=max(max_M * self.K * dtype.itemsize, 4096) * 2 * max(self.world_size, 2),
                                           rank=self.rank, world_size=self.world_size, local_world_size=self.world_size)

    def _init_gemm_ar_ctx(self, max_M, dtype=torch.bfloat16):
        N, K = self.wo.shape
        self.gemm_ar_ctx = GemmARLayer(self.group, max_M, N, K, dtype, dtype, self.world_size, persistent=True,
                                       use_ll_kernel=max_M <= 256, copy_to_local=False)

    def finalize(self):
        for c in (self.ag_ctx, self.rs_ctx, self.ar_ctx, self.gemm_ar_ctx):
            if c is not None:
                c.finalize()
        self.ag_ctx = self.rs_ctx = self.ar_ctx = self.gemm_ar_ctx = None

    # ------------------------------------------------------------------ attention core
    @torch.inference_mode()
    def _attn_core(self, qkv: torch.Tensor, bsz: int, q_len: int, position_ids: torch.Tensor, cos_sin_cache: torch.Tensor,
                   kv_cache, layer_idx: int) -> torch.Tensor:
        """qkv [bsz*q_len, q_size + 2*kv_size] -> attention output [bsz*q_len, q_size]."""
        k_cache, v_cache, _ = kv_cache.update_kv_cache(None, None, layer_idx)  # [B, S, Hkv, D]
        pos = position_ids.reshape(-1).to(torch.int32)
        bidx = torch.arange(bsz, device=qkv.device, dtype=torch.int32).repeat_interleave(q_len)
        if self.bqkv is not None:
            qkv = qkv + self.bqkv
        q = qk_norm_rope_update_kvcache(qkv.contiguous(), self.q_norm_w, self.k_norm_w, cos_sin_cache, pos, k_cache, v_cache,
                                        self.num_q_heads, self.num_kv_heads, self.head_dim, self.q_norm_eps, batch_idx=bidx)
        scale = 1.0 / math.sqrt(self.head_dim)
        if q_len == 1:
            kv_lens = pos + 1
            out = dense_decode(q, k_cache, v_cache, kv_lens, scale)  # [B, Hq, D]
            return out.view(bsz, self.q_size)
        # prefill: causal attention over the cache prefix [0, last_pos]
        if (q.is_cuda and q.dtype == torch.bfloat16 and self.head_dim == 128 and k_cache.is_contiguous()
                and os.environ.get("TDIST_ATTN_SDPA", "0") != "1"):
            # own tcgen05 flash attention (csrc/kernels/sp_attn_sm100.cu) straight on the dense cache: the queries of a
            # sequence are the last q_len of its kv_len positions (fresh prompt or chunked prefill); no host read
            from ...kernels.nvidia.sp_ag_attention_intra_node import flash_attn_varlen

            S = k_cache.shape[1]
            lens = (position_ids[:, -1] + 1).to(torch.int32)
            cu_k = torch.zeros(bsz + 1, dtype=torch.int32, device=q.device)
            cu_k[1:] = torch.cumsum(lens, 0)
            cu_q = torch.arange(bsz + 1, dtype=torch.int32, device=q.device) * q_len
            out = flash_attn_varlen(q.view(bsz * q_len, self.num_q_heads, self.head_dim), k_cache.view(-1, self.num_kv_heads, self.head_dim),
                                    v_cache.view(-1, self.num_kv_heads, self.head_dim), cu_q, cu_k, q_len, causal=True, k_batch_stride=S)
            return out.view(bsz * q_len, self.q_size)
        kv_len = int(position_ids.max().item()) + 1
        g = self.num_q_heads // self.num_kv_heads
        qh = q.view(bsz, q_len, self.num_q_heads, self.head_dim).transpose(1, 2)
        kh = k_cache[:bsz, :kv_len].transpose(1, 2)
        vh = v_cache[:bsz, :kv_len].transpose(1, 2)
        if kv_len == q_len:
            out = F.scaled_dot_product_attention(qh, kh, vh, is_causal=True, enable_gqa=g > 1)
        else:
            kpos = torch.arange(kv_len, device=q.device)[None, None, :]
            mask = (kpos <= position_ids[:, :, None]).unsqueeze(1)  # [B, 1, q_len, kv_len]
            out = F.scaled_dot_product_attention(qh, kh, vh, attn_mask=mask, enable_gqa=g > 1)
        return out.transpose(1, 2).reshape(bsz * q_len, self.q_size)

    # ------------------------------------------------------------------ forwards
    @torch.inference_mode()
    def torch_fwd(self, x, position_ids, cos_sin_cache, kv_cache, layer_idx: int):
        bsz, q_len, _ = x.shape
        qkv = F.linear(x.view(bsz * q_len, -1), self.wqkv)
        out = self._attn_core(qkv, bsz, q_len, position_ids, cos_sin_cache, kv_cache, layer_idx)
        out = F.linear(out, self.wo).view(bsz, q_len, -1)
        if self.world_size > 1:
            dist.all_reduce(out, dist.ReduceOp.SUM, group=self.group)
        return out

    @torch.inference_mode()
    def dist_triton_fwd(self, x, position_ids, cos_sin_cache, kv_cache, layer_idx: int, ag_gemm_persistent=True,
                        gemm_rs_persistent=True, autotune=False):
        """x is batch-sharded [B/W, q_len, H]; so is the output (Megatron-SP style)."""
        bsz_loc, q_len, d = x.shape
        bsz = bsz_loc * self.world_size
        qkv = ag_gemm(x.reshape(-1, d), self.wqkv, ctx=self.ag_ctx)
        out = self._attn_core(qkv, bsz, q_len, position_ids, cos_sin_cache, kv_cache, layer_idx)
        out = gemm_rs(out.contiguous(), self.wo, self.rs_ctx)
        return out.view(bsz_loc, q_len, -1)

    @torch.inference_mode()
    def dist_triton_AR_fwd(self, x, position_ids, cos_sin_cache, kv_cache, layer_idx: int):
        bsz, q_len, _ = x.shape
        qkv = linear(x.view(bsz * q_len, -1), self.wqkv)
        out = self._attn_core(qkv, bsz, q_len, position_ids, cos_sin_cache, kv_cache, layer_idx)
        out = linear(out.contiguous(), self.wo)
        if self.world_size > 1:
            out = all_reduce(out.contiguous(), method=self.ar_method, ctx=self.ar_ctx)
        return out.view(bsz, q_len, -1)

    @torch.inference_mode()
    def dist_triton_gemm_ar_fwd(self, x, position_ids, cos_sin_cache, kv_cache, layer_idx: int):
        bsz, q_len, _ = x.shape
        qkv = linear(x.view(bsz * q_len, -1), self.wqkv)
        out = self._attn_core(qkv, bsz, q_len, position_ids, cos_sin_cache, kv_cache, layer_idx)
        out = self.gemm_ar_ctx.forward(out.contiguous(), self.wo)
        return out.view(bsz, q_len, -1)
