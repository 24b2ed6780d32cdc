"""The pieces of the MoE overlap schedules that do not need a second GPU, on one device (or on the CPU paths):

  * the stage-gated AG + grouped GEMM as rank ``r`` of a pretend world W would run it — all shards are already local, so
    what is checked is the per-stage block masks, the stage-filtered gather kernel and that a grouped GEMM launch leaves
    the rows of masked-out blocks untouched — against the one-shot gather + grouped GEMM;
  * the column-chunked grouped GEMM + reduce-scatter pipeline on two streams at world size 1 against the one-launch form.

    python -m triton_dist.test.nvidia.test_moe_overlap_single_gpu
(the multi-rank forms: test_ag_moe --overlap, test_moe_reduce_rs --overlap)"""
import dataclasses

import torch

from triton_dist.kernels.nvidia.moe_reduce_rs import create_moe_rs_context, run_moe_reduce_rs
from triton_dist.kernels.nvidia.moe_utils import (grouped_gemm, moe_align_sort, moe_gather_rows, moe_gather_rows_stage)
from triton_dist.kernels.nvidia.threadblock_swizzle_ag_moe import threadblock_swizzle_ag_moe
from triton_dist.test.common import all_ok, base_parser, check_close, finish, rand, setup
from triton_dist.utils import dist_print


def staged_as_rank(full_x, w, ids, num_experts, r, W):
    """What _ag_group_gemm_staged computes on rank r once the shards are there (no transfer)."""
    m_loc = full_x.shape[0] // W
    sort = moe_align_sort(ids, num_experts)
    sched = threadblock_swizzle_ag_moe(sort, m_loc, r, W)
    n_blocks = sort.block_expert.numel()
    stage_of_block = torch.zeros(n_blocks, dtype=torch.int32, device=full_x.device)
    n_sched = min(n_blocks, sched.segment_end.numel())
    stage_of_block[:n_sched] = sched.segment_end[:n_sched]
    a_sorted = torch.full((sort.max_rows, full_x.shape[1]), 7.0, dtype=full_x.dtype, device=full_x.device)  # poison
    y = torch.full((sort.max_rows, w.shape[1]), -3.0, dtype=full_x.dtype, device=full_x.device)
    skip = torch.full_like(sort.block_expert, -1)
    seen = torch.zeros(n_blocks, dtype=torch.bool, device=full_x.device)
    for stage in range(W):
        mask = (stage_of_block == stage) & (sort.block_expert >= 0)
        moe_gather_rows_stage(full_x, sort, stage_of_block, stage, a_sorted)
        grouped_gemm(a_sorted, w, sort, out=y, block_expert=torch.where(stage_of_block == stage, sort.block_expert, skip))
        assert not bool((seen & mask).any()), "a block was computed in two stages"
        seen |= mask
    assert bool((seen == (sort.block_expert >= 0)).all()), "a block was never computed"
    return y, sort


if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=1024, help="global tokens")
    p.add_argument("--hidden", type=int, default=512)
    p.add_argument("--inter", type=int, default=256)
    p.add_argument("-E", type=int, default=8)
    p.add_argument("--topk", type=int, default=2)
    p.add_argument("--pretend-world", type=int, default=4)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = torch.bfloat16 if dev == "cuda" else torch.float32
    ok = True
    # ---- stage-gated AG + grouped GEMM, seen from every rank of the pretend world
    torch.manual_seed(3)
    x = rand((args.M, args.hidden), dtype)
    w1 = rand((args.E, args.inter, args.hidden), dtype) * 0.1
    ids = torch.rand(args.M, args.E, device=dev).topk(args.topk, dim=-1).indices.int()
    sort0 = moe_align_sort(ids, args.E)
    ref = grouped_gemm(moe_gather_rows(x, sort0), w1, sort0)
    rows = int(sort0.meta[0])
    valid = torch.repeat_interleave(sort0.block_expert[:rows // 128] >= 0, 128)
    for r in range(args.pretend_world):
        y, sort = staged_as_rank(x, w1, ids, args.E, r, args.pretend_world)
        good = bool(torch.equal(y[:rows][valid], ref[:rows][valid]))  # same kernels on the same rows: bit-identical
        if not good:
            dist_print(f"staged schedule as rank {r}: mismatch", prefix=True)
        ok = ok and good
    # ---- column-chunked grouped GEMM + reduce-scatter on two streams (world size of this launch)
    i_loc = args.inter
    w2 = rand((args.E, 1024, i_loc), dtype) * 0.1
    wts = torch.softmax(torch.randn(args.M, args.topk, device=dev), -1)
    h_sorted = torch.zeros(sort0.max_rows, i_loc, device=dev, dtype=dtype)
    h_sorted[sort0.pair_to_row.long()] = rand((args.M * args.topk, i_loc), dtype)
    outs = []
    for overlap in (False, True):
        ctx = create_moe_rs_context(rank, world, world, args.M, 1024, args.E, args.topk, dtype, overlap=overlap)
        for it in range(3):  # several calls: the scatter workspace alternates halves, the streams re-join every call
            o = run_moe_reduce_rs(h_sorted, w2, sort0, wts, ctx)
        outs.append(o.float())
        ctx.finalize()
    ok = check_close(outs[1], outs[0], dtype if dev == "cuda" else torch.float16, "chunked MoE + RS vs one launch") and ok
    ok = all_ok(ok)
    finish(ok)
