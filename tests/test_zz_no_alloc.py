"""Decode attention for callers that cannot allocate device memory.

``tdist_flash_decode_noalloc`` keeps its split-KV scratch inside the library instead of taking a buffer from the caller.
That is the entry point for code that runs where allocation is off limits: inside a serving loop whose memory pool owns
the whole GPU, in a plugin that must not touch the host application's allocator, or from plain C with no allocator at
hand. This test drives it the way such a caller would — through the C ABI with raw pointers — and checks that it
returns the same attention output as ``tdist_flash_decode`` (which is given an explicit scratch buffer), twice in a row
with different inputs, because the built-in scratch is shared between calls and still holds the previous call's
partials when the second one starts.

Collected last on purpose (file name): if it ever fails, a ``-x`` run has already exercised everything else."""
import ctypes
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _decls(lib):
    vp, i32, f32, ll = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_longlong
    lib.tdist_flash_decode_noalloc.restype = i32
    lib.tdist_flash_decode_noalloc.argtypes = [i32, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, ll, ll, ll, vp, i32, i32, vp]
    lib.tdist_flash_decode.restype = i32
    lib.tdist_flash_decode.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, ll, ll, ll, vp, i32, i32, vp]
    lib.tdist_last_error.restype = ctypes.c_char_p


def _case(seed, B, Hq, Hkv, kv_len, page=16):
    torch.manual_seed(seed)
    D = 128
    pages = (kv_len + page - 1) // page
    q = torch.randn(B, Hq, D, device="cuda").bfloat16()
    kc = torch.randn(B * pages, page, Hkv, D, device="cuda").bfloat16()
    vc = torch.randn(B * pages, page, Hkv, D, device="cuda").bfloat16()
    bt = torch.randperm(B * pages, device="cuda").int().view(B, pages).contiguous()
    lens = torch.tensor([kv_len - 5 * i for i in range(B)], dtype=torch.int32, device="cuda")
    return q, kc, vc, bt, lens


def _torch_decode_fp32(q, kc, vc, bt, lens, scale):
    """Plain PyTorch fp32 paged GQA decode attention: the independent oracle for both entry points."""
    B, Hq, D = q.shape
    page, Hkv = kc.shape[1], kc.shape[2]
    out = torch.empty(B, Hq, D, device=q.device, dtype=torch.float32)
    for b in range(B):
        n = int(lens[b])
        pages = bt[b, : (n + page - 1) // page].long()
        k = kc[pages].reshape(-1, Hkv, D)[:n].float().repeat_interleave(Hq // Hkv, dim=1)  # [n, Hq, D]
        v = vc[pages].reshape(-1, Hkv, D)[:n].float().repeat_interleave(Hq // Hkv, dim=1)
        s = torch.einsum("hd,nhd->hn", q[b].float(), k) * scale
        out[b] = torch.einsum("hn,nhd->hd", torch.softmax(s, dim=-1), v)
    return out


def test_noalloc_decode_matches_buffered_decode_twice():
    from triton_dist import _lib

    lib = ctypes.CDLL(str(_lib.lib_path()))  # raw C ABI, as a non-Python caller would bind it
    _decls(lib)
    stream = torch.cuda.current_stream().cuda_stream
    splits, D = 8, 128
    scale = 1.0 / math.sqrt(D)
    # two calls in a row, different shapes: the second one starts with the first one's partials in the shared scratch
    for seed, (B, Hq, Hkv, kv_len) in enumerate([(2, 32, 8, 777), (3, 16, 2, 300)]):
        q, kc, vc, bt, lens = _case(seed, B, Hq, Hkv, kv_len)
        out_new = torch.full((B, Hq, D), float("nan"), device="cuda", dtype=torch.bfloat16)
        out_ref = torch.empty_like(out_new)
        partial = torch.empty((B, Hq, splits, D + 1), device="cuda", dtype=torch.float32)  # the buffer the old entry point needs
        common = (B, Hq, Hkv, D, splits, scale, kc.stride(0), kc.stride(1), kc.stride(2), bt.data_ptr(), kc.shape[1], bt.shape[1], stream)
        rc = lib.tdist_flash_decode(0, q.data_ptr(), kc.data_ptr(), vc.data_ptr(), lens.data_ptr(), partial.data_ptr(),
                                    out_ref.data_ptr(), None, *common)
        assert rc == 0, lib.tdist_last_error()
        rc = lib.tdist_flash_decode_noalloc(0, q.data_ptr(), kc.data_ptr(), vc.data_ptr(), lens.data_ptr(), out_new.data_ptr(), *common)
        assert rc == 0, lib.tdist_last_error()
        torch.cuda.synchronize()
        torch.testing.assert_close(out_new.float(), out_ref.float(), atol=2e-2, rtol=2e-2)
        # and both against plain fp32 PyTorch, so that a fault shared by the two entry points cannot hide
        torch.testing.assert_close(out_new.float(), _torch_decode_fp32(q, kc, vc, bt, lens, scale), atol=2e-2, rtol=2e-2)


def test_noalloc_decode_rejects_oversized_problem():
    from triton_dist import _lib

    lib = ctypes.CDLL(str(_lib.lib_path()))
    _decls(lib)
    q, kc, vc, bt, lens = _case(0, 1, 64, 8, 64)
    out = torch.empty((1, 64, 128), device="cuda", dtype=torch.bfloat16)
    # 1 * 64 heads * 128 splits = 8192 partial rows > the 4096 built in
    rc = lib.tdist_flash_decode_noalloc(0, q.data_ptr(), kc.data_ptr(), vc.data_ptr(), lens.data_ptr(), out.data_ptr(), 1, 64, 8, 128,
                                        128, 0.1, kc.stride(0), kc.stride(1), kc.stride(2), bt.data_ptr(), kc.shape[1], bt.shape[1],
                                        torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    assert b"exceeds" in lib.tdist_last_error()
