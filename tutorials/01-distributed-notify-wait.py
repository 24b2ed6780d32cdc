"""Tutorial 1 — symmetric tensors, peer views and signals (the native equivalent of the reference's
tutorials/01-distributed-notify-wait.py).  Run:  bash scripts/launch.sh tutorials/01-distributed-notify-wait.py

Every rank allocates the SAME sequence of symmetric tensors; `symm_peer_view(t, p)` aliases rank p's copy through the
VMM mapping, so a plain PyTorch copy into it is a one-sided put over NVLink. Signals are 32-bit flags in the same heap."""
import torch
import torch.distributed as dist

from triton_dist.kernels.common_ops import CMP_GE, set_signal, wait_eq
from triton_dist.utils import dist_print, finalize_distributed, initialize_distributed, symm_peer_view, symm_tensor

initialize_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
data = symm_tensor((1024,), torch.float32)
flag = symm_tensor((world,), torch.int32)
nxt, prv = (rank + 1) % world, (rank - 1) % world
if torch.cuda.is_available():
    symm_peer_view(data, nxt).fill_(float(rank))                     # put: write into my right neighbour's tensor
    set_signal(symm_peer_view(flag, nxt)[rank:rank + 1], 1)          # notify it (release ordered after the payload)
    wait_eq(flag[prv:prv + 1], 1, cmp=CMP_GE)                        # my stream waits for my left neighbour's signal
    torch.cuda.synchronize()
    dist_print(f"rank {rank} received {data[0].item()} from rank {prv}", allowed_ranks="all")
else:
    dist_print("needs GPUs: on CPU the package falls back to torch.distributed collectives")
finalize_distributed()
