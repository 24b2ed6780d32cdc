"""1-GPU probe: per-task timeline of a TP8-shaped chain of megakernel linears (where does the time between ops go?)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.mega_triton_kernel import ModelBuilder

dev, dt = "cuda", torch.bfloat16
H, I, QKV, O = 5120, 3200, 1280, 1024
b = ModelBuilder(0, 1, 1)
x = torch.randn(1, H, device=dev, dtype=dt)
nw = torch.ones(H, device=dev, dtype=dt)
bufs = []
cur = x
for layer in range(3):
    wq = torch.randn(QKV, H, device=dev, dtype=dt) * 0.02
    wo = torch.randn(H, O, device=dev, dtype=dt) * 0.02
    w1 = torch.randn(2 * I, H, device=dev, dtype=dt) * 0.02
    w2 = torch.randn(H, I, device=dev, dtype=dt) * 0.02
    q = torch.zeros(1, QKV, device=dev, dtype=dt)
    o = torch.zeros(1, H, device=dev, dtype=dt)
    a = torch.zeros(1, I, device=dev, dtype=dt)
    d = torch.zeros(1, H, device=dev, dtype=dt)
    b.make_linear(cur, wq, q, rms_weight=nw)
    b.make_linear(q[:, :O].contiguous() if False else q[:, :O], wo, o) if False else b.make_linear(torch.zeros(1, O, device=dev, dtype=dt), wo, o)
    b.make_linear(o, w1, a, rms_weight=nw, fuse_silu=True)
    b.make_linear(a, w2, d)
    bufs += [wq, wo, w1, w2, q, o, a, d]
    cur = d
b.compile()
for _ in range(3):
    b.run()
torch.cuda.synchronize()
b.enable_profiling(True)
b.run()
torch.cuda.synchronize()
ev = b.get_trace()
t0 = min(e[5] for e in ev if e[5] > 0)
print("info", b.info())
for cta in (0, 73, 147):
    row = [(e[1], e[2], (e[5] - t0) / 1e3, (e[6] - t0) / 1e3, [(x - e[5]) / 1e3 for x in e[7]]) for e in ev if e[0] == cta]
    print(f"cta {cta}: " + "  ".join(f"op{r[1]}[{r[2]:.1f}-{r[3]:.1f} | +{r[4][0]:.1f} +{r[4][1]:.1f} +{r[4][2]:.1f}]" for r in row))
# per op: first start, last end over CTAs
ops = {}
for e in ev:
    o = ops.setdefault(e[2], [1e18, 0, 1e18, 0])
    o[0] = min(o[0], e[5]); o[1] = max(o[1], e[6]); o[2] = min(o[2], e[6]); o[3] = max(o[3], e[5])
for k in sorted(ops):
    o = ops[k]
    print(f"op {k}: first start {(o[0]-t0)/1e3:.1f} last start {(o[3]-t0)/1e3:.1f} first end {(o[2]-t0)/1e3:.1f} last end {(o[1]-t0)/1e3:.1f} us")
