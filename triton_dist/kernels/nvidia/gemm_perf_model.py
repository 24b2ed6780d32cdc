"""Analytic GEMM roofline (reference ``kernels/nvidia/gemm_perf_model.py:97-236``), with B200 entries and the
driver-measured peaks of this pool as the default denominators."""
from __future__ import annotations

import torch

from ...utils import measured_peaks

# dense tensor-core peaks in TFLOP/s (nominal) and HBM GB/s
DEVICE_TABLE = {
    "NVIDIA B200": {"bf16": 2250.0, "fp16": 2250.0, "fp8": 4500.0, "fp4": 9000.0, "tf32": 1125.0, "dram_gbs": 7700.0, "sms": 148},
    "NVIDIA H100": {"bf16": 989.0, "fp16": 989.0, "fp8": 1979.0, "tf32": 494.0, "dram_gbs": 3350.0, "sms": 132},
    "NVIDIA H800": {"bf16": 989.0, "fp16": 989.0, "fp8": 1979.0, "tf32": 494.0, "dram_gbs": 3350.0, "sms": 132},
    "NVIDIA H20": {"bf16": 148.0, "fp16": 148.0, "fp8": 296.0, "tf32": 74.0, "dram_gbs": 4000.0, "sms": 78},
    "NVIDIA A100": {"bf16": 312.0, "fp16": 312.0, "tf32": 156.0, "dram_gbs": 2039.0, "sms": 108},
}


def _device_entry(name: str = None):
    if name is None:
        name = torch.cuda.get_device_name(0) if torch.cuda.is_available() else "NVIDIA B200"
    for k, v in DEVICE_TABLE.items():
        if k.split()[-1] in name:
            return v
    return DEVICE_TABLE["NVIDIA B200"]


def get_tensorcore_tflops(dtype=torch.bfloat16, device_name: str = None, measured: bool = True) -> float:
    """Peak dense TFLOP/s; ``measured`` returns the cuBLAS burst figure of MEASURED_PEAKS.json for bf16/fp16."""
    if measured and dtype in (torch.bfloat16, torch.float16) and device_name is None:
        return measured_peaks()["bf16_tflops"]
    key = {torch.bfloat16: "bf16", torch.float16: "fp16", torch.float32: "tf32"}.get(dtype, "fp8")
    return _device_entry(device_name)[key]


def get_dram_gbps(device_name: str = None, measured: bool = True) -> float:
    return measured_peaks()["hbm_gbs"] if measured and device_name is None else _device_entry(device_name)["dram_gbs"]


def estimate_gemm_sol_time_ms(M: int, N: int, K: int, dtype=torch.bfloat16, device_name: str = None) -> float:
    """Speed-of-light time: max(FLOPs / tensor peak, bytes / DRAM bandwidth)."""
    flops = 2.0 * M * N * K
    esize = torch.empty((), dtype=dtype).element_size()
    nbytes = (M * K + N * K + M * N) * esize
    return max(flops / (get_tensorcore_tflops(dtype, device_name) * 1e9), nbytes / (get_dram_gbps(device_name) * 1e6))
