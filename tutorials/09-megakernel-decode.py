"""Tutorial 9 — one-launch-per-token decoding with the megakernel (reference docs/mega_triton_kernel.md).
Run:  RANDOM_PARAMS=1 bash scripts/launch.sh tutorials/09-megakernel-decode.py"""
import torch
import torch.distributed as dist

from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
from triton_dist.models import ModelConfig
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
cfg = ModelConfig(model_name="Qwen/Qwen3-8B", max_length=2048, dtype=torch.bfloat16, rank=rank, world_size=world)
model = Qwen3Model(batch_size=1, model_config=cfg, builder=ModelBuilder(rank, world, world))
model.set_kv_len(0)
tok = torch.tensor([1225], device=model.input_ids.device)
out = []
for _ in range(16):
    model.mega_forwrad(tok)          # (sic) one cooperative kernel launch = one full decode step
    tok = model.next_ids.clone()
    out.append(int(tok))
dist_print("generated token ids:", out, "| graph:", model.builder.info())
model.finalize()
finalize_distributed()
