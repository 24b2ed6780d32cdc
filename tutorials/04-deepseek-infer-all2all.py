"""Tutorial 4 — expert-parallel inference step with token dispatch / combine (reference
tutorials/04-deepseek-infer-all2all.py).  Run:  bash scripts/launch.sh tutorials/04-deepseek-infer-all2all.py

Each rank owns E/W experts. dispatch() sends every (token, k) pair to the owner of its expert (optionally as fp8 with
1x128 scales); the experts run as one grouped GEMM on the received rows; combine() brings the outputs back and applies
the router weights."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.moe_utils import grouped_gemm, moe_align_sort, moe_gather_rows
from triton_dist.layers.nvidia import EPAll2AllLayer
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
dev = "cuda" if torch.cuda.is_available() else "cpu"
dtype = torch.bfloat16 if dev == "cuda" else torch.float32
T, H, E, topk = 256, 1024, 8 * world, 4
x = torch.randn(T, H, device=dev, dtype=dtype)
probs = torch.softmax(torch.randn(T, E, device=dev), -1)
w, ids = probs.topk(topk, dim=-1)
layer = EPAll2AllLayer(None, T, H, topk, rank, E, world, world, dtype=dtype)
buf, counts, n = layer.dispatch(x, ids.int(), exact=False)       # rows grouped by local expert; counts on the device
epr = E // world
w_exp = torch.randn(epr, H, H, device=dev, dtype=dtype) * 0.02   # one square "expert" matrix each
n_recv = int(n.item())
local_expert = torch.repeat_interleave(torch.arange(epr, device=dev), counts.long())[:, None]
sort = moe_align_sort(local_expert.int(), epr)                   # topk = 1: one "pair" per received row
y_sorted = grouped_gemm(moe_gather_rows(buf[:n_recv].to(dtype), sort), w_exp, sort)
y = y_sorted[sort.pair_to_row.long()]                            # back to dispatch order
out = layer.combine(y, w)                                        # [T, H]
dist_print(f"rank {rank}: received {n_recv} rows, output {tuple(out.shape)}", allowed_ranks="all")
layer.finalize()
finalize_distributed()
