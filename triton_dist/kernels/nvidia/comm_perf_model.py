"""Analytic communication model (reference ``kernels/nvidia/comm_perf_model.py:94-116``) for one NVSwitch node:
every GPU reaches every peer at full NVLink bandwidth, so collective time is bytes-per-GPU / per-direction bandwidth."""
from __future__ import annotations

from ...utils import NVLINK_GBS_MEASURED


def get_nvlink_gbps(measured: bool = True) -> float:
    return NVLINK_GBS_MEASURED if measured else 900.0


def estimate_all_gather_time_ms(nbytes_per_rank: int, world_size: int, gbps: float = None) -> float:
    gbps = gbps or get_nvlink_gbps()
    return (world_size - 1) * nbytes_per_rank / (gbps * 1e6)


def estimate_reduce_scatter_time_ms(nbytes_total: int, world_size: int, gbps: float = None) -> float:
    gbps = gbps or get_nvlink_gbps()
    return (world_size - 1) / world_size * nbytes_total / (gbps * 1e6)


def estimate_all_reduce_time_ms(nbytes: int, world_size: int, gbps: float = None, nvls: bool = True) -> float:
    """NVLS: each GPU sends its data once and receives the reduced result once; ring/two-shot P2P: 2(W-1)/W."""
    gbps = gbps or get_nvlink_gbps()
    factor = 1.0 if nvls else 2.0 * (world_size - 1) / world_size
    return factor * nbytes / (gbps * 1e6)


def fused_op_roofline_ms(flops_per_gpu: float, link_bytes_per_gpu: float, tflops: float, gbps: float = None) -> float:
    """Roofline of a fused compute+collective kernel: the slower of compute at peak and bytes over NVLink."""
    gbps = gbps or get_nvlink_gbps()
    return max(flops_per_gpu / (tflops * 1e9), link_bytes_per_gpu / (gbps * 1e6))
