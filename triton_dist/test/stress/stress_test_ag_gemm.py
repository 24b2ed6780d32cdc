"""Stress: random shapes every iteration, then repeated launches to provoke hangs (mirrors reference
test/stress/stress_test_ag_gemm.py:81-130). Every iteration is checked against the fp32 oracle."""
import random

import torch

from triton_dist.kernels.nvidia.allgather_gemm import ag_gemm, create_ag_gemm_context, torch_ag_gemm
from triton_dist.kernels.nvidia.gemm_reduce_scatter import create_gemm_rs_context, gemm_rs, torch_gemm_rs
from triton_dist.test.common import all_ok, base_parser, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--rounds", type=int, default=20)
    p.add_argument("--max-m", type=int, default=4096)
    args = p.parse_args()
    group, rank, world = setup()
    dtype = torch.bfloat16 if torch.cuda.is_available() else torch.float32
    K, N = 1024, 1024
    rng = random.Random(1234)  # same shapes on every rank
    B = rand((N // world * world // world, K), dtype, 0.1)
    B2 = rand((K, N // world), dtype, 0.1)
    ag = create_ag_gemm_context(rand((args.max_m // world, K), dtype), B, rank, world, max_M=args.max_m, num_local_ranks=world)
    rs = create_gemm_rs_context(args.max_m, K, rank, world, world, dtype)
    ok = True
    for it in range(args.rounds):
        m_loc = rng.randrange(8, args.max_m // world // 8 + 1) * 8
        A = rand((m_loc, K), dtype)
        skew(True)
        C = ag_gemm(A, B, ag)
        good = check_close(C, torch_ag_gemm(A.float(), B.float()), dtype if dtype != torch.float32 else torch.float16, f"ag_gemm round {it} m_loc={m_loc}")
        H = rand((m_loc * world, N // world), dtype)
        skew(True)
        Y = gemm_rs(H, B2, rs)
        good = good and check_close(Y, torch_gemm_rs(H.float(), B2.float()), dtype if dtype != torch.float32 else torch.float16, f"gemm_rs round {it}")
        ok = all_ok(good)
        if not ok:
            break
    if ok:  # verify_hang: many back-to-back launches without host sync
        A = rand((args.max_m // world, K), dtype)
        for _ in range(50):
            ag_gemm(A, B, ag)
    ag.finalize()
    rs.finalize()
    finish(ok)
