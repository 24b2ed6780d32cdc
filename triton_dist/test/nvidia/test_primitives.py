"""Primitive-level checks (mirrors reference test_ring_put.py, test_notify.py, test_common_ops.py): peer views of
symmetric tensors, ring put with signals, notify SET/ADD, stream barrier under random skew, multimem reproducibility."""
import torch
import torch.distributed as dist

from triton_dist.kernels.common_ops import CMP_GE, BarrierAllContext, barrier_all_on_stream, set_signal, wait_eq
from triton_dist.runtime import get_heap
from triton_dist.test.common import all_ok, base_parser, finish, setup, skew
from triton_dist.utils import dist_print, nvshmem_create_tensors, symm_tensor

if __name__ == "__main__":
    p = base_parser(__doc__)
    args = p.parse_args()
    group, rank, world = setup()
    if not torch.cuda.is_available():
        dist_print("primitive tests need the CUDA symmetric heap: skipped")
        finish(True)
    heap = get_heap()
    ok = True
    # ---- ring put: write my rank into the next rank's buffer, signal, wait for the previous rank's value
    data = nvshmem_create_tensors((1024,), torch.int32, rank, world)
    sig = symm_tensor((world,), torch.int32)
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    for it in range(1, 6):
        skew(args.stress)
        data[nxt].fill_(rank * 100 + it)  # store to the peer-mapped view = put
        set_signal(heap.peer_view(sig, nxt)[rank:rank + 1], it)
        wait_eq(sig[prv:prv + 1], it, cmp=CMP_GE)
        good = bool((data[rank] == prv * 100 + it).all())
        ok = ok and good
        barrier_all_on_stream()
    ok = all_ok(ok)
    dist_print(f"ring put: {'ok' if ok else 'FAILED'}")
    # ---- notify ADD: every rank adds 1 to rank 0's counter
    ctr = symm_tensor((1,), torch.int32)
    barrier_all_on_stream()
    set_signal(heap.peer_view(ctr, 0), 1, add=True)
    barrier_all_on_stream()
    torch.cuda.synchronize()
    good = rank != 0 or int(ctr.item()) == world
    ok = all_ok(ok and good)
    dist_print(f"notify add: {'ok' if good else 'FAILED'}")
    # ---- barrier under skew: a value written before the barrier is visible to every peer after it
    bctx = BarrierAllContext(True)
    box = symm_tensor((world,), torch.int32)
    for it in range(1, 9):
        skew(True)
        for peer in range(world):
            heap.peer_view(box, peer)[rank:rank + 1].fill_(it)
        barrier_all_on_stream(bctx)
        good = bool((box == it).all())
        ok = ok and good
        barrier_all_on_stream(bctx)
    ok = all_ok(ok)
    dist_print(f"barrier stress: {'ok' if ok else 'FAILED'}")
    # ---- multimem.ld_reduce is run-to-run reproducible (reference test_multimem_ld_reduce.py:104-160)
    if heap.multicast and world > 1:
        from triton_dist.kernels.allreduce import AllReduceMethod
        from triton_dist.kernels.nvidia.allreduce import all_reduce, create_allreduce_ctx

        actx = create_allreduce_ctx(8 * 2**20, rank, world, world)
        torch.manual_seed(rank)
        x = torch.randn(1 << 18, device="cuda").bfloat16()
        y0 = all_reduce(x, method=AllReduceMethod.OneShot_Multimem, ctx=actx).clone()
        same = all(torch.equal(all_reduce(x, method=AllReduceMethod.OneShot_Multimem, ctx=actx), y0) for _ in range(5))
        ok = all_ok(ok and same)
        dist_print(f"multimem ld_reduce reproducible across runs: {same}")
        actx.finalize()
    finish(ok)
