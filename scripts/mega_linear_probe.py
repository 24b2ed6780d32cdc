"""1-GPU probe: HBM bandwidth reached by the megakernel's LINEAR task on decode-shaped weight matrices."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.mega_triton_kernel import ModelBuilder

dev = "cuda"
dt = torch.bfloat16
L = 6
for (N, K, pair, tag) in [(10240, 5120, False, "qkv W=1"), (5120, 8192, False, "o W=1"), (25600, 5120, True, "gate_up W=1"),
                          (5120, 25600, False, "down W=1"), (1280, 5120, False, "qkv TP8"), (5120, 1024, False, "o TP8"),
                          (3200, 5120, True, "gate_up TP8"), (5120, 3200, False, "down TP8")]:
    ws = [torch.randn((2 * N if pair else N), K, device=dev, dtype=dt) * 0.02 for _ in range(L)]
    x = torch.randn(1, K, device=dev, dtype=dt)
    ys = [torch.zeros(1, N, device=dev, dtype=dt) for _ in range(L)]
    b = ModelBuilder(0, 1, 1)
    for w, y in zip(ws, ys):
        b.make_linear(x, w, y, fuse_silu=pair)  # chained by default: each waits for the previous op
    b.compile()
    for _ in range(3):
        b.run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        b.run()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    t = ts[len(ts) // 2]
    byts = sum(w.numel() * 2 for w in ws)
    ref = x.float() @ ws[0].float().t()
    if pair:
        ref = torch.nn.functional.silu(ref[:, :N].to(dt).float()) * ref[:, N:].to(dt).float()
    err = (ys[0].float() - ref).abs().max().item()
    print(f"{tag:12s} N={N} K={K} pair={pair}: {t * 1e3 / L:.1f} us/op  {byts / t / 1e9:.2f} TB/s  err {err:.3g}  {b.info()}", flush=True)
    b.finalize()
