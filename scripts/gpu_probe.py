"""Developer probe for a 1-GPU box: heap bring-up, GEMM numerics, GEMM throughput vs cuBLAS.
Writes gpurun_out/probe.log style text to stdout."""
import os
import sys
import time
import traceback
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

os.environ.setdefault("RANK", "0")
os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("LOCAL_RANK", "0")
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29578")


def section(name):
    print(f"\n===== {name} =====", flush=True)


def timeit(fn, iters=20, warmup=5, flush=None):
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
    return ts[len(ts) // 2], ts[0]


def main():
    section("device")
    print(torch.cuda.get_device_name(0), torch.cuda.get_device_capability(0), torch.cuda.device_count(), "GPUs")
    from triton_dist import _lib
    from triton_dist.kernels.nvidia import gemm_persistent

    section("gemm numerics")
    for (M, N, K, bn) in [(128, 256, 64, 256), (128, 256, 256, 256), (256, 512, 1024, 128), (512, 1344, 1024, 224),
                          (1000, 1032, 520, 0), (8192, 1792, 4096, 0)]:
        try:
            torch.manual_seed(0)
            a = (torch.rand(M, K, device="cuda") * 2 - 1).bfloat16()
            b = ((torch.rand(N, K, device="cuda") * 2 - 1) * 0.1).bfloat16()
            c = gemm_persistent(a, b, block_n=bn)
            torch.cuda.synchronize()
            ref = a.float() @ b.float().t()
            err = (c.float() - ref).abs().max().item()
            print(f"M={M} N={N} K={K} bn={bn}: max abs err {err:.4g} (ref max {ref.abs().max().item():.3g})", flush=True)
        except Exception:
            traceback.print_exc()
            return

    section("gemm perf vs cuBLAS (median / best ms)")
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")
    for (M, N, K) in [(8192, 14336, 4096), (8192, 1792, 4096), (8192, 4096, 1792), (8192, 8192, 8192), (4096, 4096, 4096)]:
        a = (torch.rand(M, K, device="cuda") * 2 - 1).bfloat16()
        b = ((torch.rand(N, K, device="cuda") * 2 - 1) * 0.1).bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        med, best = timeit(lambda: torch.matmul(a, b.t(), out=out), flush=flush)
        print(f"M={M} N={N} K={K} cuBLAS      : {med:.4f} / {best:.4f} ms  {fl / med / 1e9:.0f} TFLOP/s", flush=True)
        for bn in (0, 256, 224, 192, 128):
            try:
                med, best = timeit(lambda: gemm_persistent(a, b, out=out, block_n=bn), flush=flush)
                print(f"M={M} N={N} K={K} tdist bn={bn:3d}: {med:.4f} / {best:.4f} ms  {fl / med / 1e9:.0f} TFLOP/s", flush=True)
            except Exception:
                traceback.print_exc()

    section("heap (world=1)")
    try:
        from triton_dist.utils import initialize_distributed, finalize_distributed
        from triton_dist.runtime import get_heap
        initialize_distributed(seed=0)
        h = get_heap()
        print("heap mode", h.mode, "bytes", h.nbytes, "multicast", h.multicast, flush=True)
        from triton_dist.kernels.nvidia.allreduce import all_reduce, create_allreduce_ctx
        from triton_dist.kernels.allreduce import AllReduceMethod
        ctx = create_allreduce_ctx(64 * 2**20, 0, 1, 1)
        x = torch.randn(1 << 20, device="cuda").bfloat16()
        for m in (AllReduceMethod.OneShot, AllReduceMethod.TwoShot, AllReduceMethod.OneShot_Multimem, AllReduceMethod.TwoShot_Multimem):
            y = all_reduce(x, method=m, ctx=ctx)
            torch.cuda.synchronize()
            print(m.name, "max err", (y.float() - x.float()).abs().max().item(), flush=True)
        finalize_distributed()
    except Exception:
        traceback.print_exc()


if __name__ == "__main__":
    main()
