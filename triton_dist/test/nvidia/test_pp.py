"""Pipeline-parallel activation hand-off with CommOp (mirrors reference test/nvidia/test_pp.py:77-120): rank r writes an
activation into its symmetric buffer, signals rank r+1, which waits on its stream and pulls the rows."""
import torch
import torch.distributed as dist

from triton_dist.layers.nvidia import CommOp
from triton_dist.test.common import all_ok, base_parser, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--tokens", type=int, default=512)
    p.add_argument("--hidden", type=int, default=1024)
    args = p.parse_args()
    group, rank, world = setup()
    if not torch.cuda.is_available() or world == 1:
        dist_print("test_pp needs >= 2 GPUs: skipped")
        finish(True)
    op = CommOp(rank, world, args.tokens, args.hidden, torch.bfloat16)
    ok = True
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    for it in range(1, args.verify_iters + 1):
        torch.manual_seed(1000 * it + rank)
        mine = rand((args.tokens, args.hidden), torch.bfloat16)
        torch.manual_seed(1000 * it + prv)
        expect = rand((args.tokens, args.hidden), torch.bfloat16)
        skew(args.stress)
        buf = it % op.num_buffers
        op.get_buffer(buf, args.tokens).copy_(mine)
        op.set_signal(nxt, buf, it)           # "my buffer `buf` holds generation `it`"
        op.wait_signal(prv, buf, it)          # stream waits until the previous stage published generation `it`
        got = op.read(prv, buf, args.tokens)
        good = torch.equal(got, expect)
        if not good:
            dist_print(f"iter {it}: received activation differs", allowed_ranks="all", prefix=True)
        ok = all_ok(good)
        if not ok:
            break
        dist.barrier()  # a buffer is reused every num_buffers iterations: consumers are done before the next write
    op.finalize()
    finish(ok)
