"""Small-message all-gather latency: every low-latency mode (LL packets, NVLS broadcast, pull) for 1 KB .. 256 KB per
rank, the regime of the distributed flash-decode partials (mirrors reference test/nvidia/test_ag_small_msg.py).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_ag_small_msg"""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.low_latency_allgather import create_fast_allgather_context, fast_allgather
from triton_dist.runtime import get_heap
from triton_dist.test.common import all_ok, base_parser, finish, setup
from triton_dist.utils import dist_print, perf_func_max_over_ranks, symm_tensor

if __name__ == "__main__":
    p = base_parser(__doc__)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    ctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=1 << 22)
    modes = ["pull", "push2d", "push2d_ll", "push2d_ll_multimem", "push_multimem"] if dev == "cuda" else ["pull"]
    ok = True
    nbytes = 1024
    while nbytes <= 256 * 1024:
        n = nbytes // 4
        line = f"{nbytes:>7d} B/rank:"
        for mode in modes:
            buf = symm_tensor((world, n), torch.int32)
            get_heap().barrier_all()  # every rank has zero-filled its new buffer before anybody pushes into it
            good = True
            for it in range(args.verify_iters):
                buf.zero_()
                buf[rank].fill_(rank * 1000 + it + 1)
                get_heap().barrier_all()  # a faster peer must not push into `buf` before the zero fill above has run
                fast_allgather(buf, ctx, mode=mode)
                if dev == "cuda":
                    torch.cuda.synchronize()
                exp = (torch.arange(world, device=buf.device, dtype=torch.int32) * 1000 + it + 1)[:, None].expand(world, n)
                good = good and bool(torch.equal(buf, exp))
                get_heap().barrier_all()  # all ranks have checked before the next iteration wipes the buffer
            ok = all_ok(good) and ok
            if dev == "cuda" and not args.check:
                _, t = perf_func_max_over_ranks(lambda: fast_allgather(buf, ctx, mode=mode), args.iters, args.warmup, use_graph=True)
                line += f"  {mode} {t * 1e3:6.1f} us"
            get_heap().free(buf)
        if dev == "cuda" and not args.check:
            flat = torch.empty(world * n, dtype=torch.int32, device=dev)
            src = torch.zeros(n, dtype=torch.int32, device=dev)
            _, tn = perf_func_max_over_ranks(lambda: dist.all_gather_into_tensor(flat, src), args.iters, args.warmup, use_graph=True)
            line += f"  | NCCL {tn * 1e3:6.1f} us"
        dist_print(line)
        nbytes *= 4
    ctx.finalize()
    finish(ok)
