"""Copy-engine all-gather with per-source signals (the producer of the reference's AG+GEMM, allgather.py:81-260) and the
kernel all-gathers, against torch (mirrors reference test/nvidia/test_all_gather.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_all_gather -M 8192 -K 4096"""
import torch
import torch.distributed as dist

from triton_dist.kernels.common_ops import CMP_GE, wait_eq
from triton_dist.kernels.nvidia.allgather import cp_engine_producer_all_gather_intra_node
from triton_dist.kernels.nvidia.low_latency_allgather import create_fast_allgather_context, fast_allgather
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, finish, rand, setup, skew
from triton_dist.utils import dist_print, nvshmem_barrier_all_on_stream, nvshmem_create_tensors, symm_tensor

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=2048, help="global rows")
    p.add_argument("-K", type=int, default=1024)
    args = p.parse_args()
    group, rank, world = setup()
    if not torch.cuda.is_available():
        dist_print("the copy-engine all-gather needs CUDA peer mappings: skipped")
        finish(True)
    dtype = DTYPES[args.dtype]
    m = args.M // world
    bufs = nvshmem_create_tensors((args.M, args.K), dtype, rank, world)  # peer views of one symmetric tensor
    sigs = nvshmem_create_tensors((world,), torch.int32, rank, world)
    ag_stream = torch.cuda.Stream()
    gctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=args.M * args.K * 2)
    ok = True
    for it in range(1, args.verify_iters + 1):
        torch.manual_seed(it * 100 + rank)
        x = rand((m, args.K), dtype)
        bufs[rank][rank * m:(rank + 1) * m].copy_(x)
        nvshmem_barrier_all_on_stream()
        ag_stream.wait_stream(torch.cuda.current_stream())
        skew(args.stress)
        cp_engine_producer_all_gather_intra_node(rank, world, x, bufs, ag_stream, sigs, signal_target=it, for_correctness=args.stress)
        for src in range(world):
            if src != rank:
                wait_eq(sigs[rank][src:src + 1], it, cmp=CMP_GE)  # consumer side: the main stream waits per source shard
        ref = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(ref, x)
        good = bool(torch.equal(bufs[rank], torch.cat(ref)))
        # kernel all-gather (NVLS push) of the same data
        sb = symm_tensor((world, m * args.K), dtype)
        sb[rank].copy_(x.reshape(-1))
        nvshmem_barrier_all_on_stream()  # every rank has zero-filled its new buffer before any peer pushes into it
        fast_allgather(sb, gctx, mode="push_multimem")
        torch.cuda.synchronize()
        good = good and bool(torch.equal(sb.view(args.M, args.K), torch.cat(ref)))
        from triton_dist.runtime import get_heap

        get_heap().free(sb)
        nvshmem_barrier_all_on_stream()
        ok = all_ok(good)
        if not ok:
            dist_print(f"iter {it}: all-gather mismatch", allowed_ranks="all", prefix=True)
            break
    if ok and not args.check:
        x = rand((m, args.K), dtype)

        def ce():
            ag_stream.wait_stream(torch.cuda.current_stream())
            cp_engine_producer_all_gather_intra_node(rank, world, x, bufs, ag_stream)
            torch.cuda.current_stream().wait_stream(ag_stream)

        t = bench(ce, args, flush_l2=False)
        flat = torch.empty(args.M, args.K, device="cuda", dtype=dtype)
        tn = bench(lambda: dist.all_gather_into_tensor(flat, x), args, flush_l2=False)
        nb = (world - 1) / world * args.M * args.K * 2
        dist_print(f"all_gather M={args.M} K={args.K} W={world}: copy engine {t * 1e3:.1f} us ({nb / t / 1e6:.0f} GB/s busbw) | "
                   f"NCCL {tn * 1e3:.1f} us ({nb / tn / 1e6:.0f} GB/s)")
    gctx.finalize()
    finish(ok)
