"""tcgen05 flash-attention forward vs torch SDPA and flash-attn 2 on one GPU (bf16, head_dim 128, causal + full).
    python -m triton_dist.benchmark.bench_flash_attn"""
import math

import torch
import torch.nn.functional as F

from triton_dist.kernels.nvidia.sp_ag_attention_intra_node import flash_attn_varlen


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


if __name__ == "__main__":
    for B, L, Hq, Hkv, causal in [(4, 4096, 32, 8, True), (1, 16384, 32, 8, True), (4, 4096, 32, 8, False), (32, 128, 64, 8, True)]:
        q = torch.randn(B * L, Hq, 128, device="cuda").bfloat16()
        k = torch.randn(B * L, Hkv, 128, device="cuda").bfloat16()
        v = torch.randn(B * L, Hkv, 128, device="cuda").bfloat16()
        cu = torch.arange(B + 1, device="cuda", dtype=torch.int32) * L
        out = torch.empty_like(q)
        t = timed(lambda: flash_attn_varlen(q, k, v, cu, cu, L, causal=causal, out=out))
        qh, kh, vh = (x.view(B, L, -1, 128).transpose(1, 2) for x in (q, k, v))
        ts = timed(lambda: F.scaled_dot_product_attention(qh, kh, vh, is_causal=causal, enable_gqa=True))
        line = ""
        try:
            from flash_attn import flash_attn_varlen_func

            tf = timed(lambda: flash_attn_varlen_func(q, k, v, cu, cu, L, L, causal=causal))
            line = f" | flash-attn2 {tf:.3f} ms"
        except Exception as ex:  # noqa: BLE001
            line = f" | flash-attn2 unavailable ({type(ex).__name__})"
        fl = 4.0 * B * L * L * Hq * 128 * (0.5 if causal else 1.0)
        print(f"B={B} L={L} Hq={Hq} Hkv={Hkv} causal={causal}: tcgen05 {t:.3f} ms ({fl / t / 1e9:.0f} TFLOP/s) | torch SDPA {ts:.3f} ms "
              f"({fl / ts / 1e9:.0f}){line}", flush=True)
