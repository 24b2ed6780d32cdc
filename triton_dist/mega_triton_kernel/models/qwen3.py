"""Qwen3 (dense) decode step as one megakernel launch (reference ``mega_triton_kernel/models/qwen3.py:83-201``,
layer builders ``models/layers/{tp_attn,tp_mlp}.py``).

``Qwen3Model`` takes the sharded weights of a ``triton_dist.models.DenseLLM`` (HF checkpoint or random init), allocates
the static activation buffers and the paged KV cache, and records the whole decode step with ``ModelBuilder``:

  embedding -> L x [ qkv_proj(+input RMSNorm) -> q/k norm + RoPE + KV append -> attention split -> combine
                     -> o_proj -> all-reduce(+residual) -> fc1(+post RMSNorm, SiLU*up) -> fc2 -> all-reduce(+residual) ]
            -> lm_head(+final RMSNorm) -> argmax

``fused=False`` records the reference's op-by-op graph instead (separate rms_norm / silu_mul_up / add / barrier ops)."""
from __future__ import annotations

from typing import Optional

import torch

from ...models.config import ModelConfig
from ...models.dense import DenseLLM
from ..core.builder import ModelBuilder
from .paged_kv_cache import PagedKVCache


class Qwen3Model:

    def __init__(self, batch_size: int, model_config: ModelConfig, builder: ModelBuilder, dense: Optional[DenseLLM] = None,
                 page_size: int = 16, fused: bool = True, group=None, num_kv_splits: int = 16):
        assert 1 <= batch_size <= 8, "the decode megakernel handles 1..8 sequences"
        self.builder = builder
        self.batch_size = batch_size
        self.dense = dense if dense is not None else DenseLLM(model_config, group)
        d = self.dense
        assert d.dtype == torch.bfloat16 or not torch.cuda.is_available(), "the megakernel is bf16"
        self.rank, self.world_size = d.rank, d.world_size
        self.hidden_size, self.head_dim, self.num_layers = d.hidden_size, d.head_dim, d.num_layers
        self.max_length = d.max_length
        self.fused = fused
        self.num_kv_splits = num_kv_splits
        dev, dt = d.embed_tokens.device, d.dtype
        a0 = d.layers[0].attn
        self.Hq, self.Hkv = a0.num_q_heads, a0.num_kv_heads
        self.kv_cache = PagedKVCache(self.num_layers, batch_size, self.max_length, self.Hkv, self.head_dim, page_size, dt,
                                     world_size=1, device=dev)
        B, H = batch_size, self.hidden_size
        inter_loc = d.layers[0].mlp.down_proj.shape[1]
        self.input_ids = torch.zeros((B,), dtype=torch.int64, device=dev)
        self.hidden = torch.zeros((B, H), dtype=dt, device=dev)  # residual stream
        self.qkv = torch.zeros((B, 1, self.Hq + 2 * self.Hkv, self.head_dim), dtype=dt, device=dev)
        self.q_rope = torch.zeros((B, 1, self.Hq, self.head_dim), dtype=dt, device=dev)
        self.attn_out = torch.zeros((B, 1, self.Hq, self.head_dim), dtype=dt, device=dev)
        self.act = torch.zeros((B, inter_loc), dtype=dt, device=dev)
        self.fc1_out = torch.zeros((B, 2 * inter_loc), dtype=dt, device=dev)
        self.normed = torch.zeros((B, H), dtype=dt, device=dev)
        self.tmp = torch.zeros((B, H), dtype=dt, device=dev)
        symm = builder.create_symm_tensor if (self.world_size > 1 or self._has_heap()) else \
            (lambda shape, dtype: torch.zeros(shape, dtype=dtype, device=dev))
        self.o_partial = symm((2, B, H), dt)
        self.down_partial = symm((2, B, H), dt)
        self.logits = torch.zeros((B, d.lm_head.shape[0]), dtype=dt, device=dev)
        self.next_ids = torch.zeros((B,), dtype=torch.int64, device=dev)
        import os

        # qk-norm + RoPE + attention + combine as one task type (TDIST_MEGA_FUSED_ATTN=1; off by default)
        self.fused_attn = os.environ.get("TDIST_MEGA_FUSED_ATTN", "0") == "1"
        self.build_fwd()

    @staticmethod
    def _has_heap() -> bool:
        from ...runtime import symm as _symm

        return _symm.heap_initialized()

    # ------------------------------------------------------------------ graph
    def build_fwd(self):
        b, d = self.builder, self.dense
        B, H = self.batch_size, self.hidden_size
        stride = B * H * self.hidden.element_size()
        kv = self.kv_cache
        b.make_embedding(self.input_ids, d.embed_tokens, self.hidden)
        for li, layer in enumerate(d.layers):
            attn, mlp = layer.attn, layer.mlp
            kc, vc = kv.layer(li)
            assert attn.bqkv is None, "qkv bias is not supported by the megakernel"
            if self.fused:
                b.make_qkv_proj(self.hidden, attn.wqkv, self.qkv.view(B, -1), li, rms_weight=layer.input_norm_w, rms_eps=layer.input_norm_eps)
            else:
                b.make_rms_norm(self.hidden, layer.input_norm_w, self.normed, layer.input_norm_eps, li)
                b.make_qkv_proj(self.normed, attn.wqkv, self.qkv.view(B, -1), li)
            if self.fused and self.fused_attn and b.supports_fused_attn(self.q_rope.shape[2], kc.shape[2]):
                b.make_attn_fused(self.qkv, kc, vc, kv.block_table, kv.positions, attn.q_norm_w, attn.k_norm_w, d.cos_sin_cache,
                                  self.attn_out, attn.q_norm_eps, layer_id=li, num_kv_splits=self.num_kv_splits)
            else:
                b.make_qk_norm_rope_update_kvcache(self.qkv, kc, vc, kv.block_table, kv.positions, attn.q_norm_w, attn.k_norm_w,
                                                   d.cos_sin_cache, self.q_rope, attn.q_norm_eps, attn.k_norm_eps, layer_id=li)
                b.make_attn(self.q_rope, kc, vc, kv.block_table, kv.positions, self.attn_out, layer_id=li,
                            num_kv_splits=self.num_kv_splits)
            if self.fused:
                b.make_o_proj(self.attn_out.view(B, -1), attn.wo, self.o_partial[0], li, parity_stride_bytes=stride)
                b.make_allreduce(self.o_partial[0], None, layer_id=li, residual=self.hidden, parity_stride_bytes=stride)
                b.make_fc1(self.hidden, mlp.gate_up_proj, self.act, li, rms_weight=layer.post_norm_w, rms_eps=layer.post_norm_eps,
                           fuse_silu=True)
                b.make_fc2(self.act, mlp.down_proj, self.down_partial[0], li, parity_stride_bytes=stride)
                b.make_allreduce(self.down_partial[0], None, layer_id=li, residual=self.hidden, parity_stride_bytes=stride)
            else:
                b.make_o_proj(self.attn_out.view(B, -1), attn.wo, self.o_partial[0], li, parity_stride_bytes=stride)
                b.make_allreduce(self.o_partial[0], self.tmp, layer_id=li, parity_stride_bytes=stride)
                b.make_add(self.hidden, self.tmp, self.hidden, li)
                b.make_rms_norm(self.hidden, layer.post_norm_w, self.normed, layer.post_norm_eps, li)
                b.make_fc1(self.normed, mlp.gate_up_proj, self.fc1_out, li)
                b.make_silu_mul_up(self.fc1_out, self.act, li)
                b.make_fc2(self.act, mlp.down_proj, self.down_partial[0], li, parity_stride_bytes=stride)
                b.make_allreduce(self.down_partial[0], self.tmp, layer_id=li, parity_stride_bytes=stride)
                b.make_add(self.hidden, self.tmp, self.hidden, li)
            # (no make_prefetch tasks here, unlike the reference's graph: the kernel's weight stream already runs ahead of
            # the dependency chain across task boundaries, see csrc/mega/megakernel.cu)
        b.make_linear(self.hidden, d.lm_head, self.logits, self.num_layers, rms_weight=d.norm_weight, rms_eps=d.norm_variance_epsilon)
        b.make_argmax(self.logits, self.next_ids, self.num_layers)
        b.compile()

    # ------------------------------------------------------------------ run
    def set_kv_len(self, kv_len: int, fill_random: bool = False):
        """Pretend ``kv_len`` tokens are already cached (benchmark set-up, like the reference's ctx=512 runs)."""
        if fill_random:
            self.kv_cache.rand_fill(kv_len)
        self.kv_cache.set_position(kv_len)

    @torch.inference_mode()
    def mega_forwrad(self, input_ids: torch.Tensor) -> torch.Tensor:  # (sic) the reference's spelling, qwen3.py:191
        """One decode step for ``input_ids`` [B] or [B, 1] at the cache's current position; returns logits [B, vocab].
        The sampled (argmax) token ids are in ``self.next_ids``. Advances the position."""
        self.input_ids.copy_(input_ids.reshape(-1))
        self.builder.run()
        self.kv_cache.inc()
        return self.logits

    mega_forward = mega_forwrad

    def finalize(self):
        self.builder.finalize()
