"""Tutorial 3 — hierarchical all-gather modes (reference tutorials/03-inter-node-allgather.py).
Run:  bash scripts/launch.sh tutorials/03-inter-node-allgather.py

The reference crosses nodes with NVSHMEM puts over InfiniBand and stages data per node / NUMA domain
(push_numa_2d, push_3d, ...). This framework targets ONE 8xB200 NVSwitch node: every GPU reaches every other GPU at
full NVLink bandwidth in one hop, so the hierarchical modes keep their names (code written against the reference runs
unchanged) and all map onto the flat one-hop push. The inter-node producer raises NotImplementedError with that reason."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia import cp_engine_producer_all_gather_inter_node, create_fast_allgather_context, fast_allgather
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, symm_tensor

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
ctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=1 << 22)
modes = ["push3d", "push_numa_2d", "push_numa_2d_ll", "push_numa_2d_ll_multinode"] if torch.cuda.is_available() else ["pull"]
for mode in modes:
    buf = symm_tensor((world, 4096), torch.int32)
    buf[rank].fill_(rank + 1)
    fast_allgather(buf, ctx, mode=mode)
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    ok = all(bool((buf[s] == s + 1).all()) for s in range(world))
    dist_print(f"mode {mode:28s}: {'ok' if ok else 'WRONG'}")
try:
    cp_engine_producer_all_gather_inter_node()
except NotImplementedError as e:
    dist_print(f"inter-node producer: {e}")
ctx.finalize()
finalize_distributed()
