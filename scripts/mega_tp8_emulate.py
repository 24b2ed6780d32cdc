"""1-GPU probe: Qwen3-32B decode megakernel with the TP=8 SHARD shapes of rank 0 (cross-rank all-reduces degenerate to
local adds). Gives the per-rank compute/synchronisation time of the TP8 decode step without needing 8 GPUs."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
from triton_dist.models import DenseLLM, ModelConfig

import os

for k, v in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0"), ("MASTER_ADDR", "127.0.0.1"), ("MASTER_PORT", "29577")):
    os.environ.setdefault(k, v)
from triton_dist.utils import initialize_distributed  # noqa: E402

initialize_distributed(seed=1)
tp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
name = sys.argv[2] if len(sys.argv) > 2 else "Qwen/Qwen3-32B"
cfg = ModelConfig(model_name=name, max_length=1024, dtype=torch.bfloat16, rank=0, world_size=tp)
dense = DenseLLM(cfg, None)
builder = ModelBuilder(0, 1, 1)
mega = Qwen3Model(1, cfg, builder, dense=dense)
mega.set_kv_len(512, fill_random=True)
ids = torch.full((1,), 1225, dtype=torch.long, device="cuda")


def step():
    mega.kv_cache.set_position(512)
    return mega.mega_forwrad(ids)


for _ in range(3):
    step()
torch.cuda.synchronize()
ts = []
for _ in range(10):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    step()
    e.record()
    torch.cuda.synchronize()
    ts.append(s.elapsed_time(e))
ts.sort()
print(f"{name} TP{tp}-shard decode on one GPU: {ts[len(ts) // 2]:.3f} ms/token  info={builder.info()}", flush=True)
builder.enable_profiling(True)
step()
torch.cuda.synchronize()
summ = builder.profile_summary()
print("span_us", round(summ["span_us"], 1))
for k, v in sorted(summ["per_type"].items(), key=lambda kv: -kv[1]["busy_us_per_cta"]):
    print(f"  {k:16s} busy/CTA {v['busy_us_per_cta']:9.1f} us  tasks {v['tasks']}")
# phase breakdown of linear tasks on CTA 0
ev = [e for e in builder.get_trace() if e[0] == 0 and e[3] == "linear"]
allev = [e for e in builder.get_trace() if e[0] == 0]
for typ in ("attn_split", "attn_combine", "qk_norm_rope_kv", "allreduce"):
    sel = [e for e in allev if e[3] == typ and e[7][0] > 0]
    if sel:
        print(f"CTA0 {typ}: {len(sel)} tasks  wait {sum(e[7][0] - e[5] for e in sel) / 1e3:.0f} us  run {sum(e[6] - e[7][0] for e in sel) / 1e3:.0f} us")
dep = sum(e[7][0] - e[5] for e in ev) / 1e3
stg = sum(e[7][1] - e[7][0] for e in ev) / 1e3
con = sum(e[7][2] - e[7][1] for e in ev) / 1e3
epi = sum(e[6] - e[7][2] for e in ev) / 1e3
print(f"CTA0 linear tasks: {len(ev)}  wait {dep:.0f} us  stage-x {stg:.0f} us  consume {con:.0f} us  epilogue+signal {epi:.0f} us")

# op-level timeline of one middle layer
t_all = builder.get_trace()
ops = {}
for e in t_all:
    o = ops.setdefault(e[2], [e[3], 1 << 62, 0, 0])
    o[1] = min(o[1], e[5]); o[2] = max(o[2], e[6]); o[3] += 1
keys = sorted(ops)
mid = len(keys) // 2
t0 = ops[keys[mid]][1]
print("op timeline (first begin -> last end, us) around the middle of the model:")
for k in keys[mid:mid + 12]:
    o = ops[k]
    print(f"  op {k:4d} {o[0]:16s} tiles {o[3]:4d}  {(o[1] - t0) / 1e3:8.1f} -> {(o[2] - t0) / 1e3:8.1f}")
print("CTA0 task phases in the middle of the model (us relative; begin | +deps +staged +consumed | end):")
mine = sorted([e for e in t_all if e[0] == 0], key=lambda e: e[1])
mid = len(mine) // 2
tb = mine[mid][5]
for e in mine[mid:mid + 9]:
    ph = [(x - e[5]) / 1e3 if x > 0 else float("nan") for x in e[7]]
    print(f"  {e[3]:16s} op {e[2]:4d} tile {e[4]:4d}: {(e[5] - tb) / 1e3:7.1f} | +{ph[0]:.1f} +{ph[1]:.1f} +{ph[2]:.1f} | {(e[6] - tb) / 1e3:7.1f}")

wc = builder.last_warp_cycles
print("linear main loop, cycles per warp summed over the step (CTA 0; wait-stage / ldmatrix+mma / issue-next / block-barrier):")
for w in range(16):
    print(f"  warp {w:2d}: " + "  ".join(f"{int(x):9d}" for x in wc[0, w]))
print("  mean over CTAs and warps (us at 1.9 GHz):", [round(float(wc[:, :, j].mean()) / 1900, 1) for j in range(4)])
# all tasks of one attention op in the middle of the model
att = [k for k in keys if ops[k][0] in ("attn_fused", "attn_split")]
if att:
    k = att[len(att) // 2]
    rows = sorted([e for e in t_all if e[2] == k], key=lambda e: e[6])
    t0 = min(e[5] for e in rows)
    print(f"tasks of op {k} ({ops[k][0]}): cta tile | begin +deps | end (us)")
    for e in rows:
        print(f"   cta {e[0]:3d} tile {e[4]:3d} | {(e[5] - t0) / 1e3:6.1f} +{(e[7][0] - e[5]) / 1e3:5.1f} | {(e[6] - t0) / 1e3:6.1f}")
