"""Low-latency all-gather: bit-exact check + latency (mirrors reference test/nvidia/test_fast_allgather.py)."""
import torch
import torch.distributed as dist

from triton_dist.kernels.nvidia.low_latency_allgather import (FAST_ALLGATHER_FUNC_DISPATCH, create_fast_allgather_context,
                                                              fast_allgather)
from triton_dist.test.common import all_ok, base_parser, bench, finish, setup, skew
from triton_dist.utils import bitwise_equal, dist_print, symm_tensor

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--mode", default="all")
    p.add_argument("--min-bytes", type=int, default=4096)
    p.add_argument("--max-bytes", type=int, default=8 * 2**20)
    args = p.parse_args()
    group, rank, world = setup()
    ctx = create_fast_allgather_context(rank, 0, world, 1, max_buffer_size=4 * args.max_bytes)
    modes = list(FAST_ALLGATHER_FUNC_DISPATCH) if args.mode == "all" else [args.mode]
    ok = True
    nbytes = args.min_bytes
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    while nbytes <= args.max_bytes and ok:
        buf = symm_tensor((world, nbytes // world // 4), torch.int32)
        for mode in modes:
            for it in range(args.verify_iters):
                shard = torch.randint(-2**31, 2**31 - 1, (nbytes // world // 4,), dtype=torch.int32, device=dev)
                ref = [torch.empty_like(shard) for _ in range(world)]
                if world > 1:
                    dist.all_gather(ref, shard)
                else:
                    ref = [shard]
                buf.fill_(-1)
                buf[rank].copy_(shard)
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
                dist.barrier()
                skew(args.stress)
                fast_allgather(buf, ctx, mode=mode)
                good = bitwise_equal(buf, torch.stack(ref))
                if not good:
                    dist_print(f"mode {mode} {nbytes} B iter {it}: mismatch", allowed_ranks="all", prefix=True)
                ok = all_ok(good)
                if not ok:
                    break
            if not ok:
                break
            if ok and not args.check:
                t = bench(lambda: fast_allgather(buf, ctx, mode=mode), args, flush_l2=False)
                dist_print(f"fast_allgather {mode:26s} {nbytes:>9d} B: {t * 1e3:8.1f} us  {nbytes / t / 1e6:7.1f} GB/s")
        nbytes *= 4
    ctx.finalize()
    finish(ok)
