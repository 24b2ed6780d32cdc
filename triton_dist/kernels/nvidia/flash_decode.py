"""GQA paged flash-decode (split-KV) and the log-sum-exp combines. API of the reference's
``kernels/nvidia/flash_decode.py`` (split :129-239, combines :307-532, persistent :586-760,
launchers :763-1132).

Tensors: q [B, Hq, D]; k_cache / v_cache [num_pages, page_size, Hkv, D]; block_table [B, max_pages] int32;
kv_lens [B] int32. The split kernel writes fp32 partials [B, Hq, splits, D+1] (normalised out | lse); the
combine merges them. ``*_intra_rank`` keeps the lse (output [B, Hq, D+1] fp32) so that distributed decode can
merge across ranks with ``kernel_inter_rank_gqa_fwd_batch_decode_combine_kv``.

Native kernels: csrc/kernels/decode.cu (flash_decode_split_kernel / flash_decode_combine_kernel). There is no
JIT, hence the reference's ``*_aot`` and ``*_persistent`` entry points are the same launch here.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from ... import _lib


def get_triton_combine_kv_algo_info(split_kv: int, v_head_dim: int, block_dv: Optional[int] = None) -> dict:
    """Kept for API parity (launch constants of the reference's combine kernel)."""
    return {"NUM_KV_SPLITS": split_kv, "BLOCK_DV": block_dv or 2 ** math.ceil(math.log2(v_head_dim)), "Lv": v_head_dim,
            "num_warps": 4, "num_stages": 2}


def torch_paged_decode(q, k_cache, v_cache, kv_lens, block_table, scale):
    """fp32 oracle (reference test_sp_decode_attn.py:80-131). Returns (out [B,Hq,D], lse [B,Hq])."""
    B, Hq, D = q.shape
    _, ps, Hkv, _ = k_cache.shape
    g = Hq // Hkv
    outs, lses = [], []
    for b in range(B):
        n = int(kv_lens[b])
        if n == 0:
            outs.append(torch.zeros(Hq, D, device=q.device))
            lses.append(torch.full((Hq,), float("-inf"), device=q.device))
            continue
        pages = block_table[b, :(n + ps - 1) // ps].long()
        k = k_cache[pages].reshape(-1, Hkv, D)[:n].float().repeat_interleave(g, dim=1)
        v = v_cache[pages].reshape(-1, Hkv, D)[:n].float().repeat_interleave(g, dim=1)
        s = torch.einsum("hd,nhd->hn", q[b].float(), k) * scale
        lses.append(torch.logsumexp(s, dim=-1))
        outs.append(torch.einsum("hn,nhd->hd", torch.softmax(s, dim=-1), v))
    return torch.stack(outs), torch.stack(lses)


def _decode(q, k_cache, v_cache, kv_lens, block_table, scale, kv_split, output_split, out16, out_lse):
    B, Hq, D = q.shape
    _, page_size, Hkv, _ = k_cache.shape
    if kv_split in (-1, None):
        # enough CTAs to keep ~4 per SM in flight (a CTA holds only 16 KB of K/V loads: 32 splits x 8 kv heads left a
        # 512K-token decode at 2.2 TB/s), but at least 256 positions per split; no host read of kv_lens: the cache
        # capacity bounds the length
        cap = block_table.shape[1] * page_size
        splits = max(4, min(256, (6 * 148) // max(1, B * Hkv), max(1, cap // 256)))
    else:
        splits = int(kv_split)
    if not (q.is_cuda and q.dtype in (torch.bfloat16, torch.float16) and D == 128):
        o, lse = torch_paged_decode(q, k_cache, v_cache, kv_lens, block_table, scale)
        if out16 is not None:
            out16.copy_(o.to(out16.dtype))
        if out_lse is not None:
            out_lse[..., :D].copy_(o)
            out_lse[..., D].copy_(lse)
        return
    if output_split is None or output_split.dtype != torch.float32:
        output_split = torch.empty((B, Hq, splits, D + 1), dtype=torch.float32, device=q.device)
    assert q.is_contiguous() and kv_lens.dtype == torch.int32 and block_table.dtype == torch.int32
    _lib.check(_lib.lib().tdist_flash_decode(
        _lib.dtype_code(q.dtype), q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), kv_lens.data_ptr(),
        output_split.data_ptr(), _lib.ptr(out16), _lib.ptr(out_lse), B, Hq, Hkv, D, splits, float(scale), k_cache.stride(0),
        k_cache.stride(1), k_cache.stride(2), block_table.data_ptr(), page_size, block_table.shape[1], _lib.stream_ptr()),
        "flash_decode")


def gqa_fwd_batch_decode(q, k_cache, v_cache, workspace, q_lens, kv_lens, block_table, scale, soft_cap=0.0,
                         output_split=None, output_combine=None, kv_split=-1):
    """Full decode attention on one rank: returns [B, Hq, D] in q.dtype."""
    assert soft_cap == 0.0, "soft_cap is not supported"
    if output_combine is None or output_combine.dtype != q.dtype:
        output_combine = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    _decode(q, k_cache, v_cache, kv_lens, block_table, scale, kv_split, output_split, output_combine, None)
    return output_combine


def gqa_fwd_batch_decode_intra_rank(q, k_cache, v_cache, workspace, q_lens, kv_lens, block_table, scale, soft_cap=0.0,
                                    output_split=None, output_combine=None, kv_split=-1):
    """Decode over this rank's KV shard keeping the lse: returns fp32 [B, Hq, D+1] (out | lse)."""
    assert soft_cap == 0.0, "soft_cap is not supported"
    B, Hq, D = q.shape
    if output_combine is None or output_combine.dtype != torch.float32:
        output_combine = torch.empty((B, Hq, D + 1), dtype=torch.float32, device=q.device)
    _decode(q, k_cache, v_cache, kv_lens, block_table, scale, kv_split, output_split, None, output_combine)
    return output_combine


def dense_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, kv_lens: torch.Tensor, scale: float,
                 kv_split: int = -1, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Decode attention over a DENSE cache [B, S, Hkv, D] (the layout of models.KV_Cache): q [B, Hq, D] -> [B, Hq, D]."""
    B, Hq, D = q.shape
    S, Hkv = k_cache.shape[1], k_cache.shape[2]
    if out is None:
        out = torch.empty((B, Hq, D), dtype=q.dtype, device=q.device)
    if not (q.is_cuda and q.dtype in (torch.bfloat16, torch.float16) and D == 128):
        bt = torch.arange(B, dtype=torch.int32, device=q.device)[:, None]
        o, _ = torch_paged_decode(q, k_cache[:B], v_cache[:B], kv_lens, bt, scale)
        out.copy_(o.to(q.dtype))
        return out
    if kv_split in (-1, None):  # enough CTAs to fill the GPU, at least ~256 positions per split
        kv_split = max(1, min(256, S // 256, max(1, (6 * 148) // max(B * Hkv, 1))))  # ~6 CTAs per SM (16 KB of loads each)
    partial = torch.empty((B, Hq, kv_split, D + 1), dtype=torch.float32, device=q.device)
    _lib.check(_lib.lib().tdist_flash_decode(
        _lib.dtype_code(q.dtype), q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), kv_lens.data_ptr(), partial.data_ptr(),
        out.data_ptr(), None, B, Hq, Hkv, D, kv_split, float(scale), k_cache.stride(0), k_cache.stride(1), k_cache.stride(2),
        None, 0, 0, _lib.stream_ptr()), "flash_decode(dense)")
    return out


def inter_rank_combine(all_ranks_partial: torch.Tensor, out: Optional[torch.Tensor] = None, dtype=torch.bfloat16) -> torch.Tensor:
    """all_ranks_partial: fp32 [W, B, Hq, D+1] (out | lse per rank) -> [B, Hq, D]."""
    W, B, Hq, D1 = all_ranks_partial.shape
    D = D1 - 1
    if out is None:
        out = torch.empty((B, Hq, D), dtype=dtype, device=all_ranks_partial.device)
    if not (all_ranks_partial.is_cuda and D == 128):
        lse = all_ranks_partial[..., D]
        w = torch.softmax(lse, dim=0).nan_to_num(0.0)
        out.copy_((all_ranks_partial[..., :D] * w[..., None]).sum(0).to(out.dtype))
        return out
    p = all_ranks_partial
    _lib.check(_lib.lib().tdist_flash_decode_combine(_lib.dtype_code(out.dtype), p.data_ptr(), out.data_ptr(), None, B, Hq, W,
                                                     p.stride(1), p.stride(2), p.stride(0), _lib.stream_ptr()),
               "inter_rank_combine")
    return out


# names exported by the reference (Triton kernel objects there, launchers here)
kernel_inter_rank_gqa_fwd_batch_decode_combine_kv = inter_rank_combine
kernel_gqa_fwd_batch_decode_split_kv_persistent = gqa_fwd_batch_decode
gqa_fwd_batch_decode_persistent = gqa_fwd_batch_decode


def gqa_fwd_batch_decode_aot(stream, q, k_cache, v_cache, workspace, q_lens, kv_lens, block_table, scale, soft_cap=0,
                             output_split=None, output_combine=None, kv_split=-1):
    return gqa_fwd_batch_decode(q, k_cache, v_cache, workspace, q_lens, kv_lens, block_table, scale, soft_cap, output_split,
                                output_combine, kv_split)


def gqa_fwd_batch_decode_intra_rank_aot(stream, q, k_cache, v_cache, workspace, q_lens, kv_lens, block_table, scale,
                                        soft_cap=0, output_split=None, output_combine=None, kv_split=-1):
    return gqa_fwd_batch_decode_intra_rank(q, k_cache, v_cache, workspace, q_lens, kv_lens, block_table, scale, soft_cap,
                                           output_split, output_combine, kv_split)


gqa_fwd_batch_decode_persistent_aot = gqa_fwd_batch_decode_aot
