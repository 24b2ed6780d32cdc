"""KV bandwidth of the split-KV GQA flash decode (csrc/kernels/decode.cu) on one GPU, dense cache, bs = 1.
    python -m triton_dist.benchmark.bench_flash_decode
The reference publishes ~2.6 TB/s local KV bandwidth at 100K tokens on H800 (README.md:203-205, middle panel)."""
import json
import math
import sys

import torch

from triton_dist.kernels.nvidia.flash_decode import dense_decode

if __name__ == "__main__":
    Hq, Hkv, D = 32, 8, 128
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")
    rows = []
    for kv_len in (4096, 16384, 65536, 131072, 524288):
        q = torch.randn(1, Hq, D, device="cuda").bfloat16()
        k = torch.randn(1, kv_len, Hkv, D, device="cuda").bfloat16()
        v = torch.randn(1, kv_len, Hkv, D, device="cuda").bfloat16()
        lens = torch.tensor([kv_len], dtype=torch.int32, device="cuda")
        out = torch.empty(1, Hq, D, device="cuda", dtype=torch.bfloat16)
        fn = lambda: dense_decode(q, k, v, lens, 1.0 / math.sqrt(D), out=out)  # noqa: E731
        for _ in range(3):
            fn()
        ts = []
        for _ in range(10):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ms = sorted(ts)[len(ts) // 2]
        kv_bytes = 2 * kv_len * Hkv * D * 2
        rows.append(dict(kv_len=kv_len, us=ms * 1e3, kv_gbs=kv_bytes / ms / 1e6))
        print(f"kv_len {kv_len:7d}: {ms * 1e3:8.1f} us  KV {kv_bytes / ms / 1e6:7.0f} GB/s", flush=True)
    if len(sys.argv) > 1:
        json.dump(rows, open(sys.argv[1], "w"), indent=1)
