"""Tutorial 4 — an expert-parallel MoE inference step with low-latency dispatch / combine (DeepSeek-style EP).
(Counterpart of the reference's tutorials/04-deepseek-infer-all2all.py, which builds the all-to-all from
``putmem_signal`` per (rank, expert) block with an optional fp8 payload + scale. Kernels here: csrc/kernels/ep_a2a.cu.)

    bash scripts/launch.sh tutorials/04-deepseek-infer-all2all.py

The problem. E experts are spread over W ranks (E / W each). Every rank has T tokens; the router picks top-k experts per
token. One MoE layer is
        dispatch:  every (token, k) pair travels to the rank that owns expert ids[token, k]
        experts:   each rank runs its experts on the rows it received
        combine:   every row travels back and out[token] = sum_k w[token, k] * y[token, k]
In decode T is small (128 tokens per rank is the reference's benchmark point), so both exchanges are pure latency:
the reference reports 137 us for dispatch on 32 x H800; this implementation measures 48.8 us on 8 x B200
(profiles/README.md).

STEP 0 — why not ``all_to_all_single``. NCCL needs the split sizes on the HOST: route -> count -> copy the counts to the
CPU -> exchange them -> copy back -> all-to-all. Two host synchronisations per layer, ~100 us before the first byte
moves, and nothing of it can live in a CUDA graph.

STEP 1 — dispatch without the host (``EPAll2AllLayer.dispatch``; kernels ``ep_counts_kernel`` + ``ep_dispatch_kernel``):
  (a) counts: one CTA histograms this rank's (token, k) pairs per expert in shared memory and stores its E counts into
      row ``rank`` of EVERY peer's [W][E] count table (plain NVLink stores through ``symm_at``, a few hundred bytes),
      then a cross-rank flag barrier: afterwards every rank holds the full W x E matrix;
  (b) offsets: from that matrix every sender computes, for each expert e, the FINAL base row of "rows for e coming from
      me" in the destination rank's receive buffer. The layout is expert-major, then source rank: all rows of local
      expert 0 (from rank 0, 1, ...), then expert 1, ... — exactly what a grouped GEMM wants, so the receiver never sorts
      or compacts. The receiver derives ``counts[e]`` and the total from the same matrix;
  (c) payload: one WARP per (token, k) pair takes the next row of its (expert, me) range (an atomic cursor) and stores
      the token straight there in the destination rank's symmetric buffer with 16-byte stores; the chosen row is kept in
      ``scatter_idx`` for the way back. A grid barrier plus ONE cross-rank barrier close the kernel: all rows of all ranks
      have landed;
  (d) optional fp8: the warp quantises the row on the fly to e4m3 with one fp32 scale per 128 values (DeepSeek's 1 x 128
      block scaling; 16 lanes x 8 values share a scale through shuffles) — half the bytes over NVLink, and the expert GEMM
      can consume the rows as they are (step 3).
  Two small launches, no host involvement, CUDA-graph capturable; flags carry a monotone sequence number (tutorial 1).
  ``exact=False`` returns the received-row count as a DEVICE scalar, so nothing synchronises unless the caller wants to
  slice on the host (this script does, for its tutorial-size expert loop).

STEP 2 — the experts. Received rows are grouped by local expert, ``counts[e]`` rows each. bf16 path: one GEMM per
expert (or the grouped tcgen05 GEMM with a per-128-row-block expert table, tutorial-size here: a loop).

STEP 3 — fp8 experts (``gemm_fp8``, csrc/kernels/gemm_fp8_sm100.cuh). The dispatched e4m3 rows and their 1 x 128 scales go
straight into ``tcgen05.mma.kind::f8f6f4``; the weights are quantised ONCE with one scale per 128 x 128 block. The scales
change every 128 values of K, so the kernel keeps a fresh TMEM accumulator per 128-wide k block and the epilogue warps
add ``scale_a[row] * scale_b[block] * partial`` into fp32 registers (two-level accumulation); 1.2-1.5 PFLOP/s on B200.

STEP 4 — combine (``EPAll2AllLayer.combine``, ``ep_combine_kernel``): the reverse trip. The expert outputs stay where
they are (in a symmetric buffer, rows in dispatch order); every rank PULLS the k rows of each of its tokens from the ranks
that computed them — row ``scatter_idx[token, k]`` on the owner of expert ``ids[token, k]`` — multiplies by the router
weight and sums in fp32. A pull needs no second count exchange and writes only local memory.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist
import torch.nn.functional as F

from triton_dist.layers.nvidia import EPAll2AllLayer
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, perf_func_max_over_ranks

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
gpu = torch.cuda.is_available()
dev = "cuda" if gpu else "cpu"
dtype = torch.bfloat16 if gpu else torch.float32
T, H, I, E, topk = 128, 1024, 512, 8 * world, 4          # tokens per rank, hidden, expert FFN width, experts, top-k
epr = E // world

# every rank generates ALL experts from the same seed and keeps its slice, so a dense oracle is available everywhere
g = torch.Generator(device=dev).manual_seed(1234)
router = (torch.randn(E, H, device=dev, generator=g) * 0.05).to(dtype)
w1_all = (torch.randn(E, 2 * I, H, device=dev, generator=g) * 0.03).to(dtype)    # gate and up projection, stacked
w2_all = (torch.randn(E, H, I, device=dev, generator=g) * 0.03).to(dtype)
w1, w2 = w1_all[rank * epr:(rank + 1) * epr], w2_all[rank * epr:(rank + 1) * epr]
torch.manual_seed(77 + rank)
x = (torch.randn(T, H, device=dev) * 0.5).to(dtype)


def route(x):
    p = torch.softmax(F.linear(x.float(), router.float()), -1)
    w, ids = p.topk(topk, dim=-1)
    return w / w.sum(-1, keepdim=True), ids.int()


def swiglu_expert(rows, e_w1, e_w2):
    gu = F.linear(rows, e_w1)
    return F.linear(F.silu(gu[:, :I]) * gu[:, I:], e_w2)


def dense_oracle(x):
    """What the layer must compute, with every expert available locally (fp32 accumulate over k)."""
    w, ids = route(x)
    out = torch.zeros(x.shape, device=dev)
    for k in range(topk):
        for e in ids[:, k].unique().tolist():
            m = ids[:, k] == e
            out[m] += w[m, k:k + 1] * swiglu_expert(x[m], w1_all[e], w2_all[e]).float()
    return out


# ---- steps 1, 2, 4: bf16 payload
layer = EPAll2AllLayer(None, T, H, topk, rank, E, world, world, dtype=dtype)


def moe_bf16(x):
    w, ids = route(x)
    buf, counts, n = layer.dispatch(x, ids, exact=False)      # buf: rows grouped by local expert; counts [E/W], n: device
    n_recv = int(n.item())                                     # only this tutorial-size expert loop needs it on the host
    rows = buf[:n_recv].to(dtype)
    y = torch.empty_like(rows)
    offs = (torch.cumsum(counts, 0) - counts).tolist()
    for e, (c, o) in enumerate(zip(counts.tolist(), offs)):
        if c:
            y[o:o + c] = swiglu_expert(rows[o:o + c], w1[e], w2[e])
    return layer.combine(y, w), n_recv


out, n_recv = moe_bf16(x)
ref = dense_oracle(x)
err = (out.float() - ref).abs().max().item()
dist_print(f"rank {rank}: received {n_recv} rows for my {epr} experts (mean would be {T * topk}); bf16 max |err| {err:.4f} "
           f"(|ref| max {ref.abs().max().item():.3f})", allowed_ranks="all")
assert err <= 0.05 * ref.abs().max().item() + 1e-3

# ---- steps 1(d) + 3: fp8 payload feeding the block-scaled fp8 GEMM (GPU only)
if gpu:
    from triton_dist.kernels.nvidia.gemm_fp8 import FP8, gemm_fp8, quant_fp8_128x128, quant_fp8_1x128

    layer8 = EPAll2AllLayer(None, T, H, topk, rank, E, world, world, dtype=dtype, fp8_dispatch=True)
    w1_8, s1 = quant_fp8_128x128(w1)                            # once, at load time: [E/W, 2I, H] e4m3 + [E/W, 2I/128, H/128]
    w2_8, s2 = quant_fp8_128x128(w2)

    def moe_fp8(x):
        w, ids = route(x)
        buf, counts, n = layer8.dispatch(x, ids, exact=False)
        n_recv = int(n.item())
        a8, sa = buf[:n_recv].view(FP8), layer8.dispatch_scales()[:n_recv]   # e4m3 rows + [n, H/128] scales, as they arrived
        y = torch.empty((n_recv, H), dtype=dtype, device=dev)
        offs = (torch.cumsum(counts, 0) - counts).tolist()
        for e, (c, o) in enumerate(zip(counts.tolist(), offs)):
            if c:
                gu = gemm_fp8(a8[o:o + c], sa[o:o + c], w1_8[e], s1[e], out_dtype=dtype)
                act8, sact = quant_fp8_1x128(F.silu(gu[:, :I]) * gu[:, I:])
                gemm_fp8(act8, sact, w2_8[e], s2[e], out=y[o:o + c])
        return layer8.combine(y, w)

    out8 = moe_fp8(x)
    err8 = (out8.float() - ref).abs().max().item()
    dist_print(f"rank {rank}: fp8 dispatch + fp8 experts max |err| {err8:.4f} (e4m3 has 3 mantissa bits: a few % of the scale)",
               allowed_ranks="all")
    assert err8 <= 0.12 * ref.abs().max().item()

    # ---- latency of the two exchanges alone (device time, max over ranks)
    w, ids = route(x)
    y_fake = torch.randn(T * topk * 2, H, device=dev).to(dtype)
    _, t_d = perf_func_max_over_ranks(lambda: layer.dispatch(x, ids, exact=False), 20, 5)
    _, t_d8 = perf_func_max_over_ranks(lambda: layer8.dispatch(x, ids, exact=False), 20, 5)
    _, t_c = perf_func_max_over_ranks(lambda: layer.combine(y_fake, w), 20, 5)
    dist_print(f"T={T}/rank H={H} top-{topk} E={E}: dispatch bf16 {t_d * 1e3:.1f} us | dispatch fp8 {t_d8 * 1e3:.1f} us | "
               f"combine {t_c * 1e3:.1f} us")
    torch.cuda.synchronize()
    layer8.finalize()
layer.finalize()
finalize_distributed()
