"""ctypes binding of libtdist.so (C ABI in include/tdist.h).

The native library is built in tree by ``build.py`` into ``triton_dist/lib``.
On a machine with a GPU every op goes through it; a missing library is a hard
error there (no silent PyTorch fallback). On a CPU-only machine the package
still imports and the torch/gloo plumbing backend is used instead.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path
from typing import Optional

import torch

MAX_RANKS = 16
MAX_CTAS = 160

DT_BF16, DT_F16, DT_F32, DT_FP8 = 0, 1, 2, 3

_DTYPE_MAP = {torch.bfloat16: DT_BF16, torch.float16: DT_F16, torch.float32: DT_F32}


def dtype_code(dtype: torch.dtype) -> int:
    try:
        return _DTYPE_MAP[dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype for native kernels: {dtype}") from None


class SymmCtx(ctypes.Structure):
    _fields_ = [
        ("base", ctypes.c_void_p * MAX_RANKS),
        ("mc_base", ctypes.c_void_p),
        ("bytes", ctypes.c_ulonglong),
        ("rank", ctypes.c_int),
        ("world", ctypes.c_int),
    ]


class GemmArgs(ctypes.Structure):
    _fields_ = [
        ("a", ctypes.c_void_p),
        ("b", ctypes.c_void_p),
        ("c", ctypes.c_void_p),
        ("M", ctypes.c_int),
        ("N", ctypes.c_int),
        ("K", ctypes.c_int),
        ("lda", ctypes.c_longlong),
        ("ldb", ctypes.c_longlong),
        ("ldc", ctypes.c_longlong),
        ("dtype", ctypes.c_int),
        ("block_n", ctypes.c_int),
        ("group_m", ctypes.c_int),
        ("num_sms", ctypes.c_int),
        ("transpose_out", ctypes.c_int),
        ("block_expert", ctypes.c_void_p),
        ("num_experts", ctypes.c_int),
        ("cta_group", ctypes.c_int),
    ]


class FusedArgs(ctypes.Structure):
    _fields_ = [
        ("ctx", SymmCtx),
        ("mode", ctypes.c_int),
        ("ws", ctypes.c_void_p),
        ("out", ctypes.c_void_p),
        ("flags", ctypes.c_void_p),
        ("flags2", ctypes.c_void_p),
        ("counters", ctypes.c_void_p),
        ("phase", ctypes.c_void_p),
        ("done", ctypes.c_void_p),
        ("num_comm_ctas", ctypes.c_int),
        ("ar_one_shot", ctypes.c_int),
        ("use_mc", ctypes.c_int),
        ("variant", ctypes.c_int),
        ("ws_bytes", ctypes.c_size_t),
        ("flags_words", ctypes.c_size_t),
    ]


class MegaOp(ctypes.Structure):
    _fields_ = [
        ("inp", ctypes.c_void_p * 6),
        ("out", ctypes.c_void_p * 3),
        ("M", ctypes.c_int),
        ("N", ctypes.c_int),
        ("K", ctypes.c_int),
        ("i", ctypes.c_int * 6),
        ("l", ctypes.c_longlong * 3),
        ("f", ctypes.c_float * 2),
    ]


_LIB: Optional[ctypes.CDLL] = None
_LIB_ERR: Optional[str] = None


def lib_path() -> Path:
    return Path(__file__).resolve().parent / "lib" / "libtdist.so"


def _declare(lib: ctypes.CDLL) -> None:
    vp, i32, u32, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint, ctypes.c_size_t
    P = ctypes.POINTER
    lib.tdist_last_error.restype = ctypes.c_char_p
    lib.tdist_last_error.argtypes = []

    def sig(name, args):
        fn = getattr(lib, name, None)
        if fn is None:  # older build of the library: leave undeclared, caller gets AttributeError
            return
        fn.restype = i32
        fn.argtypes = args

    sig("tdist_heap_create", [i32, i32, i32, sz, ctypes.c_char_p, P(vp)])
    sig("tdist_heap_create_ipc", [i32, i32, i32, sz, P(vp), vp])
    sig("tdist_heap_open_ipc", [vp, vp])
    sig("tdist_heap_mc_phase_a", [vp])
    sig("tdist_heap_mc_phase_b", [vp])
    lib.tdist_heap_base.restype = vp
    lib.tdist_heap_base.argtypes = [vp, i32]
    lib.tdist_heap_mc_base.restype = vp
    lib.tdist_heap_mc_base.argtypes = [vp]
    lib.tdist_heap_bytes.restype = sz
    lib.tdist_heap_bytes.argtypes = [vp]
    sig("tdist_heap_mode", [vp])
    sig("tdist_heap_fill_ctx", [vp, P(SymmCtx)])
    sig("tdist_heap_destroy", [vp])
    sig("tdist_stream_write_value32", [vp, vp, u32])
    sig("tdist_stream_wait_value32", [vp, vp, u32, i32])
    sig("tdist_stream_write_value64", [vp, vp, ctypes.c_ulonglong])
    sig("tdist_stream_wait_value64", [vp, vp, ctypes.c_ulonglong, i32])
    sig("tdist_barrier_all", [P(SymmCtx), vp, vp, vp, vp])
    sig("tdist_allreduce", [P(SymmCtx), i32, i32, vp, vp, sz, vp, sz, vp, vp, vp, i32, i32, vp])
    sig("tdist_allgather", [P(SymmCtx), i32, vp, vp, vp, sz, sz, vp, vp, vp, i32, vp])
    sig("tdist_reduce_scatter", [P(SymmCtx), i32, vp, vp, sz, vp, sz, vp, vp, vp, i32, vp])
    sig("tdist_slot_reduce", [i32, vp, vp, sz, i32, i32, i32, vp])
    sig("tdist_p2p_copy", [vp, vp, sz, i32, vp])
    sig("tdist_signal_set", [vp, u32, i32, vp])
    sig("tdist_signal_wait", [vp, u32, i32, vp])
    f32, ll = ctypes.c_float, ctypes.c_longlong
    sig("tdist_gemm_fp8", [vp, vp, ll, vp, vp, vp, i32, i32, i32, ll, ll, ll, i32, vp, i32, i32, vp])
    sig("tdist_gemv", [i32, vp, vp, vp, vp, i32, i32, i32, ll, ll, ll, vp])
    sig("tdist_rmsnorm", [i32, vp, vp, vp, vp, i32, i32, f32, ll, ll, vp])
    sig("tdist_silu_mul", [i32, vp, vp, ll, i32, vp])
    sig("tdist_add", [i32, vp, vp, vp, ll, vp])
    sig("tdist_qk_norm_rope_kv", [i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, ll, ll, ll, vp, i32, i32, vp])
    sig("tdist_flash_decode", [i32, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, ll, ll, ll, vp, i32, i32, vp])
    sig("tdist_flash_decode_combine", [i32, vp, vp, vp, i32, i32, i32, ll, ll, ll, vp])
    sig("tdist_embedding", [i32, vp, vp, vp, i32, i32, vp])
    sig("tdist_argmax", [i32, vp, vp, i32, i32, ll, vp])
    sig("tdist_moe_align_sort", [vp, i32, i32, i32, vp, vp, vp, vp, i32, vp])
    sig("tdist_primitives_selftest", [P(SymmCtx), vp, vp, u32, vp])
    sig("tdist_moe_tile_swizzle", [vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, i32, vp])
    sig("tdist_moe_gather_rows", [vp, vp, vp, vp, i32, i32, i32, i32, vp])
    sig("tdist_moe_gather_rows_stage", [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp])
    sig("tdist_moe_topk_reduce", [i32, vp, vp, vp, vp, i32, i32, i32, vp])
    sig("tdist_moe_reduce_rs", [P(SymmCtx), i32, vp, vp, sz, vp, vp, vp, i32, i32, i32, vp, vp, vp, i32, vp])
    sig("tdist_a2a", [P(SymmCtx), vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp])
    sig("tdist_ep_counts", [P(SymmCtx), vp, i32, i32, vp, vp, vp, vp, vp])
    sig("tdist_ep_dispatch", [P(SymmCtx), i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, vp,
                              i32, vp])
    sig("tdist_ep_combine", [P(SymmCtx), i32, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, i32, vp])
    sig("tdist_attn_fwd_tc", [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, ll, ll, i32, i32, i32, i32, i32, f32, vp])
    sig("tdist_sp_attn_fwd", [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, f32, vp])
    sig("tdist_sp_kv_scatter", [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp])
    sig("tdist_mega_create", [i32, i32, i32, P(vp)])
    sig("tdist_mega_add_op", [vp, i32, P(MegaOp), i32, i32, i32])
    sig("tdist_mega_compile", [vp, P(SymmCtx), vp])
    sig("tdist_mega_run", [vp, vp])
    sig("tdist_mega_info", [vp, P(i32 * 6)])
    sig("tdist_mega_op_tiles", [vp, i32])
    sig("tdist_mega_linear_plan", [i32, i32, i32, i32, i32, i32, P(i32 * 12)])
    sig("tdist_mega_get_queue", [vp, i32, vp, i32])
    sig("tdist_mega_set_trace", [vp, i32])
    sig("tdist_mega_get_trace", [vp, vp, sz])
    sig("tdist_mega_destroy", [vp])
    sig("tdist_gemm", [P(GemmArgs), vp])
    sig("tdist_gemm_tile_order", [i32, i32, i32, i32, i32, i32, i32, vp, i32])
    sig("tdist_gemm_fused", [P(GemmArgs), P(FusedArgs), vp])


def load(required: bool = False) -> Optional[ctypes.CDLL]:
    """Load libtdist.so once. ``required`` turns a failure into an exception."""
    global _LIB, _LIB_ERR
    if _LIB is None and _LIB_ERR is None:
        path = Path(os.environ.get("TDIST_LIB", str(lib_path())))
        try:
            _LIB = ctypes.CDLL(str(path))
            _declare(_LIB)
        except OSError as e:  # not built, or no CUDA runtime deps
            _LIB_ERR = f"cannot load {path}: {e} (run `python build.py`)"
    if _LIB is None and required:
        raise RuntimeError(_LIB_ERR)
    return _LIB


def lib() -> ctypes.CDLL:
    """The native library; raises if it is not available."""
    return load(required=True)


def check(rc: int, what: str = "tdist call") -> None:
    if rc != 0:
        msg = lib().tdist_last_error()
        raise RuntimeError(f"{what} failed: {msg.decode() if msg else 'unknown error'}")


def use_native() -> bool:
    """True when the CUDA kernels are the execution path (a GPU is present)."""
    return torch.cuda.is_available()


def stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()
