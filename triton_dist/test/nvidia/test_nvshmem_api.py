"""Symmetric-heap host API under its NVSHMEM-style names (mirrors reference test/nvidia/test_nvshmem_api.py):
create / free symmetric tensors, peer views, barrier_all on a stream, signal wait, allocator reuse."""
import torch

from triton_dist.runtime import get_heap
from triton_dist.test.common import all_ok, base_parser, finish, setup
from triton_dist.utils import (NVSHMEM_SIGNAL_DTYPE, dist_print, nvshmem_barrier_all_on_stream, nvshmem_create_tensor,
                               nvshmem_create_tensors, nvshmem_free_tensor_sync, symm_peer_view)

if __name__ == "__main__":
    p = base_parser(__doc__)
    args = p.parse_args()
    group, rank, world = setup()
    ok = True
    heap = get_heap()
    t = nvshmem_create_tensor((4, 256), torch.float32)
    ok = ok and heap.is_symmetric(t) and tuple(t.shape) == (4, 256)
    t.fill_(rank + 1)
    nvshmem_barrier_all_on_stream()
    if torch.cuda.is_available():
        nxt = (rank + 1) % world
        ok = ok and bool((symm_peer_view(t, nxt) == nxt + 1).all())  # load through the peer mapping = get
        views = nvshmem_create_tensors((64,), NVSHMEM_SIGNAL_DTYPE, rank, world)
        ok = ok and len(views) == world and (world == 1 or views[rank].data_ptr() != views[nxt].data_ptr())
        nvshmem_barrier_all_on_stream()
        nvshmem_free_tensor_sync(views)
    # allocator: freeing and re-allocating the same size returns the same offset on every rank (symmetry is preserved)
    off0 = heap.offset_of(t) if hasattr(heap, "offset_of") else None
    nvshmem_free_tensor_sync(t)
    t2 = nvshmem_create_tensor((4, 256), torch.float32)
    if off0 is not None:
        ok = ok and heap.offset_of(t2) == off0
    nvshmem_free_tensor_sync(t2)
    dist_print(f"[nvshmem api] heap mode {getattr(heap, 'mode', 'host')}, multicast {getattr(heap, 'multicast', False)}")
    finish(all_ok(ok))
