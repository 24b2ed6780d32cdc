"""Throughput of the block-scaled fp8 GEMM against the bf16 tcgen05 engine and cuBLAS bf16 on one GPU (L2 flushed).
    python -m triton_dist.benchmark.bench_gemm_fp8"""
import torch

from triton_dist.kernels.nvidia.allgather_gemm import gemm_persistent
from triton_dist.kernels.nvidia.gemm_fp8 import gemm_fp8, quant_fp8_1x128, quant_fp8_128x128
from triton_dist.utils import perf_func

if __name__ == "__main__":
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")

    def timed(fn, iters=10):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(iters):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        return sorted(ts)[len(ts) // 2]

    for M, N, K in [(8192, 7168, 2048), (8192, 4096, 7168), (4096, 2048, 7168), (8192, 14336, 4096)]:
        a = torch.randn(M, K, device="cuda")
        b = torch.randn(N, K, device="cuda") * 0.05
        a8, sa = quant_fp8_1x128(a)
        b8, sb = quant_fp8_128x128(b)
        a16, b16 = a.bfloat16(), b.bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        t8 = timed(lambda: gemm_fp8(a8, sa, b8, sb, out=out))
        t16 = timed(lambda: gemm_persistent(a16, b16, out=out))
        tcb = timed(lambda: torch.matmul(a16, b16.t(), out=out))
        fl = 2.0 * M * N * K
        print(f"M={M} N={N} K={K}: fp8 block-scaled {t8:.3f} ms ({fl / t8 / 1e9:.0f} TFLOP/s) | bf16 tcgen05 {t16:.3f} ms "
              f"({fl / t16 / 1e9:.0f}) | cuBLAS bf16 {tcb:.3f} ms ({fl / tcb / 1e9:.0f})")
