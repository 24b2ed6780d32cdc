"""Host runtime helpers (mirrors reference test/nvidia/test_utils.py): perf_func, dist_print, assert_allclose,
generate_data, group_profile (merged chrome trace), topology probes."""
import os
import tempfile

import torch

from triton_dist import utils
from triton_dist.test.common import all_ok, base_parser, finish, setup
from triton_dist.utils import assert_allclose, bitwise_equal, dist_print, generate_data, group_profile, perf_func

if __name__ == "__main__":
    p = base_parser(__doc__)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    ok = True
    a = torch.randn(256, 256, device=dev)
    out, ms = perf_func(lambda: a @ a, iters=3, warmup_iters=1)
    ok = ok and out.shape == (256, 256) and ms >= 0
    assert_allclose(a, a.clone(), rtol=0, atol=0, verbose=False)
    try:
        assert_allclose(a, a + 1, rtol=1e-3, atol=1e-3, verbose=False)
        ok = False
    except (AssertionError, RuntimeError):
        pass
    ok = ok and bitwise_equal(a, a.clone()) and not bitwise_equal(a, a + 1)
    gen = generate_data([((4, 8), torch.float32, (1.0, 0.0)), ((2, 2), torch.int8, (4, 0))])
    ts = next(gen) if hasattr(gen, "__next__") else gen
    ok = ok and tuple(ts[0].shape) == (4, 8) and ts[1].dtype == torch.int8
    with tempfile.TemporaryDirectory() as d:
        with group_profile("unit", do_prof=True, group=group, out_dir=d):
            (a @ a).sum().item()
        ok = ok and (rank != 0 or world == 1 or os.path.exists(os.path.join(d, "unit", "merged.json.gz")))
    info = dict(is_cuda=utils.is_cuda(), has_tma=utils.has_tma(), blackwell=utils.is_blackwell(),
                fullmesh_nvlink=utils.has_fullmesh_nvlink() if dev == "cuda" else None,
                multimem=utils.is_nvshmem_multimem_supported() if dev == "cuda" else None,
                numa_world=utils.get_numa_world_size(), nvlink_gbs=utils.get_intranode_max_speed() if dev == "cuda" else None)
    dist_print(f"[utils] {info}")
    dist_print("[utils] every rank prints this line", allowed_ranks="all", prefix=True, need_sync=True)
    finish(all_ok(ok))
