"""Local reduction of W stacked partial results, `out = sum_w in[w]` in ring order (reference perf_ring_reduce_tma.py times
its TMA / non-TMA Triton variants of this kernel). Single GPU.
    python -m triton_dist.test.nvidia.perf_ring_reduce_tma -M 8192 -N 4096 --splits 8"""
import torch

from triton_dist.kernels.nvidia.reduce_scatter import ring_reduce
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, setup
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=2048, help="rows of ONE split")
    p.add_argument("-N", type=int, default=4096)
    p.add_argument("--splits", type=int, default=8)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    x = torch.randn(args.splits * args.M, args.N, device=dev).to(dtype)
    out = torch.empty(args.M, args.N, device=dev, dtype=dtype)
    ok = True
    for begin in (0, args.splits // 2):
        ring_reduce(x, out, begin, args.splits)
        ref = x.float().view(args.splits, args.M, args.N).sum(0)
        ok = check_close(out, ref, dtype if dev == "cuda" else torch.float16, f"ring_reduce begin={begin}") and ok
    ok = all_ok(ok)
    if ok and not args.check and dev == "cuda":
        t = bench(lambda: ring_reduce(x, out, 0, args.splits), args)
        tr = bench(lambda: torch.sum(x.view(args.splits, args.M, args.N), 0, out=out), args)
        nbytes = (args.splits + 1) * args.M * args.N * x.element_size()
        dist_print(f"ring_reduce {args.splits} x [{args.M}, {args.N}]: {t * 1e3:.1f} us ({nbytes / t / 1e9:.2f} TB/s) | torch.sum {tr * 1e3:.1f} us")
    finish(ok)
