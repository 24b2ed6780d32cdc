"""1-GPU probe: CTA-pair GEMM kernel vs single-CTA kernel vs cuBLAS (numerics + throughput)."""
import sys
import traceback
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.kernels.nvidia import gemm_persistent


def timeit(fn, iters=15, warmup=4, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


print("== numerics (cta_group=2)", flush=True)
for (M, N, K, bn) in [(256, 256, 64, 256), (512, 512, 1024, 128), (1024, 1344, 1024, 224), (1000, 1032, 520, 0), (8192, 1792, 4096, 0),
                      (300, 136, 264, 64)]:
    try:
        torch.manual_seed(0)
        a = (torch.rand(M, K, device="cuda") * 2 - 1).bfloat16()
        b = ((torch.rand(N, K, device="cuda") * 2 - 1) * 0.1).bfloat16()
        c = gemm_persistent(a, b, block_n=bn, cta_group=2)
        torch.cuda.synchronize()
        ref = a.float() @ b.float().t()
        print(f"M={M} N={N} K={K} bn={bn}: max abs err {(c.float() - ref).abs().max().item():.4g} (ref max {ref.abs().max().item():.3g})", flush=True)
    except Exception:
        traceback.print_exc()
        sys.exit(1)

print("== perf (median ms, L2 flushed)", flush=True)
flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")
for (M, N, K) in [(8192, 14336, 4096), (8192, 4096, 14336), (8192, 1792, 4096), (8192, 4096, 1792), (8192, 8192, 8192)]:
    a = (torch.rand(M, K, device="cuda") * 2 - 1).bfloat16()
    b = ((torch.rand(N, K, device="cuda") * 2 - 1) * 0.1).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * M * N * K
    t = timeit(lambda: torch.matmul(a, b.t(), out=out), flush=flush)
    print(f"M={M} N={N} K={K} cuBLAS          : {t:.4f} ms {fl / t / 1e9:.0f} TFLOP/s", flush=True)
    for cg, bns in ((1, (0, 256)), (2, (0, 256, 224, 192, 128, 112))):
        for bn in bns:
            try:
                t = timeit(lambda: gemm_persistent(a, b, out=out, block_n=bn, cta_group=cg), flush=flush)
                print(f"M={M} N={N} K={K} tdist cg={cg} bn={bn:3d}: {t:.4f} ms {fl / t / 1e9:.0f} TFLOP/s", flush=True)
            except Exception:
                traceback.print_exc()

print("== gemv bandwidth", flush=True)
from triton_dist.kernels.nvidia.elementwise import gemv  # noqa: E402

for (N, K) in [(25600, 5120), (5120, 25600), (151936, 5120), (8192, 8192)]:
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.02
    for M in (1, 4):
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        y = gemv(x, w)
        ref = x.float() @ w.float().t()
        err = (y.float() - ref).abs().max().item()
        t = timeit(lambda: gemv(x, w), flush=flush)
        print(f"gemv M={M} N={N} K={K}: {t * 1e3:.1f} us {N * K * 2 / t / 1e9:.2f} TB/s err {err:.3g}", flush=True)
