"""Op-type registry (reference core/registry.py maps task classes to generated Triton code; the native interpreter has a
fixed switch, so the registry is the table of op codes and the C++ body that implements each one)."""
from __future__ import annotations

from typing import Dict, NamedTuple


class OpInfo(NamedTuple):
    code: int
    name: str
    native_body: str  # function in csrc/mega/megakernel.cu
    reference: str    # reference file the body corresponds to


_OPS = [
    OpInfo(0, "linear", "linear_body", "kernels/linear.py, kernels/mlp_fc1.py (tasks/linear.py)"),
    OpInfo(1, "rms_norm", "rmsnorm_body", "kernels/norm.py (tasks/norm.py)"),
    OpInfo(2, "silu_mul", "silu_mul_body", "kernels/activation.py (tasks/activation.py)"),
    OpInfo(3, "add", "add_body", "kernels/elementwise.py (tasks/elementwise.py)"),
    OpInfo(4, "qk_norm_rope_kv", "qk_rope_body", "kernels/norm.py:rmsnorm_rope_update_kv_cache (tasks/norm.py)"),
    OpInfo(5, "attn_split", "attn_split_body", "kernels/flash_attn.py:33 (tasks/attn.py)"),
    OpInfo(6, "attn_combine", "attn_combine_body", "kernels/flash_attn.py:166 (tasks/attn.py)"),
    OpInfo(7, "allreduce", "allreduce_body", "kernels/allreduce.py (tasks/allreduce.py)"),
    OpInfo(8, "barrier", "barrier_all_block", "kernels/barrier.py (tasks/barrier.py)"),
    OpInfo(9, "prefetch", "prefetch_body", "kernels/prefetch.py (tasks/prefetch.py)"),
    OpInfo(10, "embedding", "embedding_body", "(torch op in the reference model)"),
    OpInfo(11, "argmax", "argmax_body", "(torch op in the reference model)"),
    OpInfo(12, "attn_fused", "attn_split_body<G, true>", "kernels/norm.py + kernels/flash_attn.py:33,166 in one task"),
]

REGISTRY: Dict[str, OpInfo] = {o.name: o for o in _OPS}
BY_CODE: Dict[int, OpInfo] = {o.code: o for o in _OPS}


def op_name(code: int) -> str:
    return BY_CODE[code].name
