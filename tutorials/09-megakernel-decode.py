"""Tutorial 9 — a whole decode step as ONE persistent kernel (the "megakernel").
(Counterpart of the reference's docs/mega_triton_kernel.md and mega_triton_kernel/test: there a ``ModelBuilder`` records
tasks and GENERATES the source of one Triton kernel per model; here the same builder API feeds a native scheduler
(csrc/mega/builder.cc) and ONE pre-compiled interpreter kernel (csrc/mega/megakernel.cu) — nothing is compiled per model.
The numbering differs from the reference, whose tutorials 09 / 10 are the AMD versions of 07 / 08.)

    RANDOM_PARAMS=1 bash scripts/launch.sh tutorials/09-megakernel-decode.py

WHY. A decode step of Qwen3-32B at batch 1 is ~450 small kernels (per layer: norm, qkv GEMV, q/k norm + RoPE + KV append,
attention, combine, o_proj, all-reduce, norm, gate/up, SiLU, down, all-reduce). Each reads a few MB and runs for
5-30 us; launch gaps, the ramp-up of every kernel's weight stream and the tails between kernels cost more than the
work. A CUDA graph removes the host launches but not the ramp-up and the tails. The megakernel keeps every SM busy with ONE
resident CTA that walks a static task list; the weight stream of the next task is already in flight while the current
one finishes.

STEP 1 — record a graph. ``ModelBuilder.make_*`` calls append ops (same names as the reference: ``make_rms_norm``,
``make_fc1``, ``make_silu_mul_up``, ``make_fc2``, ``make_add``, ``make_attn``, ``make_allreduce``, ...). Tensors are ordinary
torch tensors (symmetric ones for all-reduce inputs); an op depends on the op that was recorded before it unless told
otherwise. B200 additions are optional arguments: ``rms_weight=`` folds the RMSNorm into the linear's prologue,
``fuse_silu=True`` makes fc1 emit silu(gate) * up, ``residual=`` folds the residual add into the all-reduce.

STEP 2 — ``compile()``: the native builder cuts every op into per-tile TASKS (a linear op: column tiles sized so that
all SMs get work; attention: (kv head, split) pairs; element-wise ops: row blocks), deals them round-robin into one
static queue per SM, and lays out the tables the kernel interprets: {queue, op table, TMA descriptors of the weights,
weight-stream schedule}. ``info()`` shows ops / tasks / queue length.

STEP 3 — ``run()``: one cooperative launch, one CTA per SM. Every CTA loops over its queue:
        wait   until scoreboard[producer op] >= step * tiles(producer)        (ld.acquire.gpu spin, bounded)
        run    the task body (linear: TMA-staged weight stream -> mma.sync; attention; norm; all-reduce via NVLS ...)
        signal scoreboard[my op] += 1                                          (release)
The scoreboard counters are MONOTONE — after step s an op's counter is s * tiles — so nothing is reset between tokens
and the launch can sit in a CUDA graph. Cross-rank ops (all-reduce, barrier) use the flag protocol of tutorial 1 on
symmetric buffers, with ``multimem.ld_reduce`` doing the sum in the switch.

STEP 4 — look at the timeline: ``enable_profiling()`` makes every task stamp ``globaltimer`` at its start and end;
``profile_summary()`` aggregates per op type, ``dump_trace(path)`` writes a chrome trace with one row per SM. That is how
the open problem of this kernel was found: at TP8 the streaming phases run at ~84 % of the HBM rate, but 1.7 of 4 ms per
token are dependency waits between whole ops (profiles/README.md R2-7, docs/roadmap.md).

STEP 5 — the real model: ``Qwen3Model`` records all layers (embedding ... LM head, argmax), ``mega_forwrad`` (sic, the
reference's spelling) runs one step. Measured (Qwen3-32B, bs=1, ctx=512, TP8 on 8 x B200): 5.25 ms per token against
8.88 ms for the same model through ``Engine`` with torch kernels in a CUDA graph and 6.31 ms with the fused all-reduce
backend (the reference on 8 x H800: 10.8 / 9.2 / 7.41 ms).
"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))  # runnable from a checkout without installing

import torch
import torch.distributed as dist
import torch.nn.functional as F

from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
from triton_dist.models import ModelConfig
from triton_dist.utils import (dist_print, finalize_distributed, initialize_distributed, nvshmem_create_tensor,
                               nvshmem_free_tensor_sync)

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
gpu = torch.cuda.is_available()
dev = "cuda" if gpu else "cpu"
dtype = torch.bfloat16 if gpu else torch.float32


def rms(x, w, eps=1e-6):
    xf = x.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype).float() * w.float()


# ---- steps 1-3: one tensor-parallel MLP block by hand: norm -> gate/up (SwiGLU) -> down -> all-reduce (+ residual)
torch.manual_seed(5)                                         # same weights on every rank; each keeps its slice
M, H, I = 2, 1024, 3072                                      # I / world must stay a multiple of 64 (the linear task's K)
x = torch.randn(M, H, device=dev).to(dtype)
nw = (torch.rand(H, device=dev) + 0.5).to(dtype)
w1_full = (torch.randn(2, I, H, device=dev) * H ** -0.5).to(dtype)      # [gate | up]
w2_full = (torch.randn(H, I, device=dev) * I ** -0.5).to(dtype)
Il = I // world
w1 = w1_full[:, rank * Il:(rank + 1) * Il].reshape(2 * Il, H).contiguous()   # column-parallel: my slice of gate and of up
w2 = w2_full[:, rank * Il:(rank + 1) * Il].contiguous()                      # row-parallel: my slice of the reduction dim
act = torch.zeros(M, Il, device=dev, dtype=dtype)
# all-reduce input: symmetric (peers read it through the switch) and DOUBLE-buffered by step parity — a rank that is
# already in step s + 1 writes the other half while a slower peer may still be reading step s (no barrier between steps)
partial = nvshmem_create_tensor((2, M, H), dtype)
stride = M * H * partial.element_size()
summed = torch.zeros(M, H, device=dev, dtype=dtype)
out = torch.zeros(M, H, device=dev, dtype=dtype)

bld = ModelBuilder(rank, world, world)
bld.make_fc1(x, w1, act, rms_weight=nw, fuse_silu=True)      # RMSNorm in the prologue, silu(gate) * up in the epilogue
bld.make_fc2(act, w2, partial[0], parity_stride_bytes=stride)             # depends on fc1 (the op recorded before it)
bld.make_allreduce(partial[0], summed, parity_stride_bytes=stride)         # summed = sum over ranks of the partials
bld.make_add(summed, x, out)                                              # residual (``residual=`` would fold it in, in place)
bld.compile()
dist_print(f"hand-built TP-MLP graph: {bld.info()}")

xn = rms(x, nw).to(dtype).float()
g = (xn @ w1_full[0].float().t()).to(dtype).float()
u = (xn @ w1_full[1].float().t()).to(dtype).float()
ref = ((F.silu(g) * u).to(dtype).float() @ w2_full.float().t()).to(dtype).float() + x.float()
for step in range(3):                                        # several launches: the scoreboard is never reset
    bld.run()
    if gpu:
        torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err <= 3e-2 * ref.abs().max().item(), f"step {step}: {err}"
dist_print(f"3 launches, max |err| vs the unsharded fp32 MLP: {err:.4f} (|ref| max {ref.abs().max().item():.2f})")

# ---- step 4: the per-task timeline
if gpu:
    bld.enable_profiling()
    bld.run()
    torch.cuda.synchronize()
    dist_print(f"timeline summary (us per op type): {bld.profile_summary()}")
    if rank == 0 and os.environ.get("TDIST_TUTORIAL_TRACE"):
        bld.dump_trace(os.environ["TDIST_TUTORIAL_TRACE"])   # open in chrome://tracing or Perfetto: one row per SM
bld.finalize()
nvshmem_free_tensor_sync(partial)

# ---- step 5: the whole model, one launch per token
cfg = ModelConfig(model_name=os.environ.get("TDIST_TUTORIAL_MODEL", "Qwen/Qwen3-8B"), max_length=2048, dtype=torch.bfloat16,
                  rank=rank, world_size=world)
model = Qwen3Model(batch_size=1, model_config=cfg, builder=ModelBuilder(rank, world, world))
model.set_kv_len(0)
tok = torch.tensor([1225], device=model.input_ids.device)
generated = []
if gpu:
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
for _ in range(16):
    model.mega_forwrad(tok)                                  # embedding ... all layers ... LM head ... argmax: ONE kernel
    tok = model.next_ids.clone()
    generated.append(int(tok))
if gpu:
    t1.record()
    torch.cuda.synchronize()
    dist_print(f"{cfg.model_name}, TP{world}: {t0.elapsed_time(t1) / 16:.2f} ms per token including the host loop")
dist_print("generated token ids (random-init weights unless the checkpoint is cached):", generated, "| graph:", model.builder.info())
model.finalize()
finalize_distributed()
