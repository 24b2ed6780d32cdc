"""Bandwidth-bound decode-path ops: RMSNorm(+residual), SiLU*up, add, per-head q/k norm + RoPE +
KV-cache append, weight-streaming GEMV, embedding gather, greedy sampling.

The reference takes these from flashinfer / cuBLAS inside its layers
(layers/nvidia/tp_attn.py:29-43,67,173; tp_mlp.py:209-228) and re-implements them as Triton task
bodies for the megakernel (mega_triton_kernel/kernels/{norm,activation,elementwise,linear}.py).
Here they are native CUDA kernels (csrc/kernels/decode.cu) used by both the layers and the megakernel.
Every function has a plain-torch path for CPU-only runs (the oracle in the tests).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from ... import _lib


def _native(t: torch.Tensor) -> bool:
    return t.is_cuda and t.dtype in (torch.bfloat16, torch.float16)


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6, residual: Optional[torch.Tensor] = None,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out = rmsnorm(x + residual) * weight; ``residual`` (if given) is updated in place to x + residual."""
    hidden = x.shape[-1]
    x2 = x.reshape(-1, hidden)
    if out is None:
        out = torch.empty_like(x2)
    if not _native(x):
        xs = x2.float()
        if residual is not None:
            r2 = residual.view(-1, hidden)
            r2.copy_((xs + r2.float()).to(x.dtype))
            xs = r2.float()
        out.copy_((xs * torch.rsqrt(xs.pow(2).mean(-1, keepdim=True) + eps) * weight.float()).to(x.dtype))
        return out.view(x.shape)
    assert x2.stride(1) == 1 and (residual is None or residual.is_contiguous())
    _lib.check(_lib.lib().tdist_rmsnorm(_lib.dtype_code(x.dtype), x2.data_ptr(), _lib.ptr(residual), weight.data_ptr(),
                                        out.data_ptr(), x2.shape[0], hidden, eps, x2.stride(0), out.stride(0),
                                        _lib.stream_ptr()), "rmsnorm")
    return out.view(x.shape)


def silu_mul(gate_up: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """silu(gate_up[..., :I]) * gate_up[..., I:]."""
    inter = gate_up.shape[-1] // 2
    g2 = gate_up.reshape(-1, 2 * inter)
    if out is None:
        out = torch.empty((g2.shape[0], inter), dtype=gate_up.dtype, device=gate_up.device)
    if not _native(gate_up):
        out.copy_((F.silu(g2[:, :inter].float()) * g2[:, inter:].float()).to(gate_up.dtype))
    else:
        assert g2.is_contiguous()
        _lib.check(_lib.lib().tdist_silu_mul(_lib.dtype_code(g2.dtype), g2.data_ptr(), out.data_ptr(), g2.shape[0], inter,
                                             _lib.stream_ptr()), "silu_mul")
    return out.view(*gate_up.shape[:-1], inter)


def add(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if out is None:
        out = torch.empty_like(a)
    if not _native(a) or a.numel() % 8 != 0 or not (a.is_contiguous() and b.is_contiguous()):
        torch.add(a, b, out=out)
    else:
        _lib.check(_lib.lib().tdist_add(_lib.dtype_code(a.dtype), a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(),
                                        _lib.stream_ptr()), "add")
    return out


def make_cos_sin_cache(head_dim: int, max_pos: int, theta: float = 1e6, device=None) -> torch.Tensor:
    """fp32 [max_pos, head_dim]: cos(head_dim/2) | sin(head_dim/2), NeoX pairing (i, i + head_dim/2)."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float64) / head_dim))
    ang = torch.arange(max_pos, dtype=torch.float64)[:, None] * inv[None, :]
    return torch.cat([ang.cos(), ang.sin()], dim=-1).float().to(device or ("cuda" if torch.cuda.is_available() else "cpu"))


def torch_qk_norm_rope(qkv, q_norm_w, k_norm_w, cos_sin, positions, Hq, Hkv, head_dim, eps):
    """Reference: returns (q [T,Hq,D], k [T,Hkv,D], v [T,Hkv,D]) after per-head norm and RoPE."""
    T = qkv.shape[0]
    x = qkv.view(T, Hq + 2 * Hkv, head_dim).float()
    q, k, v = x[:, :Hq], x[:, Hq:Hq + Hkv], x[:, Hq + Hkv:]

    def norm(t, w):
        if w is None:
            return t
        return t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + eps) * w.float()

    def rope(t):
        half = head_dim // 2
        cs = cos_sin[positions.long()]
        cos, sin = cs[:, None, :half], cs[:, None, half:]
        t1, t2 = t[..., :half], t[..., half:]
        return torch.cat([t1 * cos - t2 * sin, t2 * cos + t1 * sin], dim=-1)

    return rope(norm(q, q_norm_w)).to(qkv.dtype), rope(norm(k, k_norm_w)).to(qkv.dtype), v.to(qkv.dtype)


def qk_norm_rope_update_kvcache(qkv: torch.Tensor, q_norm_w: Optional[torch.Tensor], k_norm_w: Optional[torch.Tensor],
                                cos_sin: torch.Tensor, positions: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                                Hq: int, Hkv: int, head_dim: int = 128, eps: float = 1e-6,
                                batch_idx: Optional[torch.Tensor] = None, block_table: Optional[torch.Tensor] = None,
                                q_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """qkv [T, (Hq+2Hkv)*D] -> q_out [T, Hq, D]; k/v of token t are written at ``positions[t]`` of sequence
    ``batch_idx[t]`` into the cache. Dense cache: [B, S, Hkv, D]; paged: [pages, page_size, Hkv, D] + block_table."""
    T = qkv.shape[0]
    if q_out is None:
        q_out = torch.empty((T, Hq, head_dim), dtype=qkv.dtype, device=qkv.device)
    if not _native(qkv) or head_dim != 128:
        q, k, v = torch_qk_norm_rope(qkv, q_norm_w, k_norm_w, cos_sin, positions, Hq, Hkv, head_dim, eps)
        q_out.copy_(q)
        b = batch_idx.long() if batch_idx is not None else torch.zeros(T, dtype=torch.long, device=qkv.device)
        pos = positions.long()
        if block_table is None:
            k_cache[b, pos] = k
            v_cache[b, pos] = v
        else:
            ps = k_cache.shape[1]
            page = block_table[b, pos // ps].long()
            k_cache[page, pos % ps] = k
            v_cache[page, pos % ps] = v
        return q_out
    assert qkv.is_contiguous() and positions.dtype == torch.int32
    page_size = k_cache.shape[1] if block_table is not None else 0
    _lib.check(_lib.lib().tdist_qk_norm_rope_kv(
        _lib.dtype_code(qkv.dtype), qkv.data_ptr(), _lib.ptr(q_norm_w), _lib.ptr(k_norm_w), cos_sin.data_ptr(),
        positions.data_ptr(), _lib.ptr(batch_idx), q_out.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), T, Hq, Hkv,
        head_dim, eps, k_cache.stride(0), k_cache.stride(1), k_cache.stride(2), _lib.ptr(block_table), page_size,
        block_table.shape[1] if block_table is not None else 0, _lib.stream_ptr()), "qk_norm_rope_kv")
    return q_out


def gemv(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y[M, N] = x[M, K] @ w[N, K]^T for 1 <= M <= 16 (weight streaming at HBM speed)."""
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=x.dtype, device=x.device)
    if not _native(x):
        out.copy_(F.linear(x, w, bias))
        return out
    _lib.check(_lib.lib().tdist_gemv(_lib.dtype_code(x.dtype), x.data_ptr(), w.data_ptr(), _lib.ptr(bias), out.data_ptr(), M,
                                     N, K, x.stride(0), w.stride(0), out.stride(0), _lib.stream_ptr()), "gemv")
    return out


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """nn.Linear forward on native kernels: GEMV for <= 16 rows, tcgen05 GEMM otherwise."""
    from .allgather_gemm import gemm_persistent

    x2 = x.reshape(-1, x.shape[-1])
    M, K = x2.shape
    if not _native(x2) or K % 8 != 0 or w.shape[0] % 8 != 0:
        y = F.linear(x2, w, bias)
    elif M <= 16:
        y = gemv(x2, w, bias, out)
    else:
        y = gemm_persistent(x2, w, out=out)
        if bias is not None:
            y += bias
    return y.view(*x.shape[:-1], w.shape[0])


def embedding(ids: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    if not _native(table):
        return F.embedding(ids, table)
    flat = ids.reshape(-1).to(torch.int64).contiguous()
    out = torch.empty((flat.numel(), table.shape[1]), dtype=table.dtype, device=table.device)
    _lib.check(_lib.lib().tdist_embedding(_lib.dtype_code(table.dtype), flat.data_ptr(), table.data_ptr(), out.data_ptr(),
                                          flat.numel(), table.shape[1], _lib.stream_ptr()), "embedding")
    return out.view(*ids.shape, table.shape[1])


def argmax(logits: torch.Tensor) -> torch.Tensor:
    l2 = logits.reshape(-1, logits.shape[-1])
    if not _native(l2):
        return l2.float().argmax(-1).view(logits.shape[:-1])
    out = torch.empty((l2.shape[0],), dtype=torch.int64, device=l2.device)
    _lib.check(_lib.lib().tdist_argmax(_lib.dtype_code(l2.dtype), l2.data_ptr(), out.data_ptr(), l2.shape[0], l2.shape[1],
                                       l2.stride(0), _lib.stream_ptr()), "argmax")
    return out.view(logits.shape[:-1])
