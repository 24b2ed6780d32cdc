"""CPU checks of the op layer's torch paths (the oracles used by the GPU tests) and of pure-host logic."""
import math

import torch


def test_rmsnorm_cpu_matches_formula():
    from triton_dist.kernels.nvidia.elementwise import rmsnorm

    x = torch.randn(4, 64)
    w = torch.rand(64) + 0.5
    y = rmsnorm(x, w, 1e-6)
    ref = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-6) * w
    torch.testing.assert_close(y, ref)


def test_rope_cache_roundtrip_cpu():
    from triton_dist.kernels.nvidia.elementwise import make_cos_sin_cache, qk_norm_rope_update_kvcache

    T, Hq, Hkv, D = 3, 4, 2, 128
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D)
    cs = make_cos_sin_cache(D, 16, device="cpu")
    kc, vc = torch.zeros(1, 16, Hkv, D), torch.zeros(1, 16, Hkv, D)
    pos = torch.tensor([0, 1, 2], dtype=torch.int32)
    q = qk_norm_rope_update_kvcache(qkv, None, None, cs, pos, kc, vc, Hq, Hkv, D)
    # position 0 is the identity rotation
    torch.testing.assert_close(q[0], qkv[0].view(Hq + 2 * Hkv, D)[:Hq])
    torch.testing.assert_close(vc[0, 2], qkv[2].view(Hq + 2 * Hkv, D)[Hq + Hkv:])
    # rotation preserves the norm
    torch.testing.assert_close(q[2].norm(dim=-1), qkv[2].view(Hq + 2 * Hkv, D)[:Hq].norm(dim=-1))


def test_paged_decode_oracle_matches_sdpa():
    from triton_dist.kernels.nvidia.flash_decode import inter_rank_combine, torch_paged_decode

    B, Hq, Hkv, D, page, n = 2, 4, 2, 128, 4, 10
    q = torch.randn(B, Hq, D)
    kc, vc = torch.randn(8, page, Hkv, D), torch.randn(8, page, Hkv, D)
    bt = torch.tensor([[0, 1, 2, 3], [4, 5, 6, 7]], dtype=torch.int32)
    lens = torch.tensor([n, n], dtype=torch.int32)
    out, lse = torch_paged_decode(q, kc, vc, lens, bt, 1 / math.sqrt(D))
    k = kc.view(2, 16, Hkv, D)[:, :n].repeat_interleave(2, dim=2).transpose(1, 2)
    v = vc.view(2, 16, Hkv, D)[:, :n].repeat_interleave(2, dim=2).transpose(1, 2)
    ref = torch.nn.functional.scaled_dot_product_attention(q[:, :, None], k, v)[:, :, 0]
    torch.testing.assert_close(out, ref, atol=1e-5, rtol=1e-5)
    # splitting the KV in two and merging by lse is exact
    o0, l0 = torch_paged_decode(q, kc, vc, torch.tensor([4, 4], dtype=torch.int32), bt, 1 / math.sqrt(D))
    kc2, vc2 = kc.view(2, 16, Hkv, D)[:, 4:].reshape(-1, page, Hkv, D), vc.view(2, 16, Hkv, D)[:, 4:].reshape(-1, page, Hkv, D)
    bt2 = torch.tensor([[0, 1, 2], [3, 4, 5]], dtype=torch.int32)
    o1, l1 = torch_paged_decode(q, kc2, vc2, torch.tensor([6, 6], dtype=torch.int32), bt2, 1 / math.sqrt(D))
    parts = torch.stack([torch.cat([o0, l0[..., None]], -1), torch.cat([o1, l1[..., None]], -1)])
    torch.testing.assert_close(inter_rank_combine(parts, dtype=torch.float32), ref, atol=1e-5, rtol=1e-5)


def test_allreduce_method_names():
    from triton_dist.kernels.allreduce import AllReduceMethod, get_allreduce_methods, to_allreduce_method

    assert to_allreduce_method("two_shot_multimem") == AllReduceMethod.TwoShot_Multimem
    assert "one_shot" in get_allreduce_methods()


def test_tile_order_modules():
    """AG+GEMM / GEMM+RS tile orders and the AG+MoE block schedule (reference K45, K59, K60)."""
    import torch

    from triton_dist.kernels.nvidia.ag_gemm_threadblock_swizzle import ag_gemm_tile_order
    from triton_dist.kernels.nvidia.gemm_rs_threadblock_swizzle import gemm_rs_tile_order
    from triton_dist.kernels.nvidia.moe_utils import torch_align_sort
    from triton_dist.kernels.nvidia.threadblock_swizzle_ag_moe import threadblock_swizzle_ag_moe

    ids = torch.rand(512, 16).topk(4, dim=-1).indices
    s = torch_align_sort(ids, 16)
    sch = threadblock_swizzle_ag_moe(s, 128, 1, 4)
    nb = int(s.meta[1])
    assert sorted(sch.block_order[:nb].tolist()) == list(range(nb))
    e = sch.segment_end[sch.block_order[:nb].long()]
    assert bool((e[1:] >= e[:-1]).all())
    assert int(sch.stage_count.sum()) == nb
    o = ag_gemm_tile_order(1024, 1024, 4, 1)
    assert len(o) == len(set(o)) == (1024 // 128) * (1024 // 256) and o[0][0] == 1
    o = gemm_rs_tile_order(1024, 1024, 4, 1)
    assert o[0][0] == 2 and o[-1][0] == 1


def test_compile_aot_table():
    """The built library carries the kernel instantiation table and the C ABI (reference test_compile_aot.py)."""
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, "-m", "triton_dist.test.nvidia.test_compile_aot"], cwd=root, capture_output=True, text=True,
                       env={**__import__("os").environ, "PYTHONPATH": str(root)})
    assert r.returncode == 0, r.stdout + r.stderr


def test_language_primitive_map():
    """Every primitive name of the reference's language layer maps to an existing native device function."""
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, "-m", "triton_dist.test.nvidia.test_language_extra"], cwd=root, capture_output=True, text=True,
                       env={**__import__("os").environ, "PYTHONPATH": str(root)})
    assert r.returncode == 0, r.stdout + r.stderr


def test_megakernel_linear_plan_invariants():
    """Host-side geometry of the megakernel's weight stream (csrc/mega/mega.h linear_geom, builder.cc tiles_of) over the
    decode shapes of the supported models at TP 1..8: every column and every k is covered exactly by the planned stages,
    a stage fits its buffer, and at least two buffers fit next to the scratch area."""
    import ctypes

    from triton_dist import _lib

    lib = _lib.load()
    if lib is None or not hasattr(lib, "tdist_mega_linear_plan"):
        import pytest

        pytest.skip("native library not built")
    shapes = []
    for hidden, inter, heads, kv, vocab in [(5120, 25600, 64, 8, 151936), (4096, 12288, 32, 8, 151936), (1024, 3072, 16, 8, 151936),
                                            (8192, 28672, 64, 8, 128256), (5120, 27648, 80, 8, 155136)]:
        for tp in (1, 2, 4, 8):
            shapes += [((heads + 2 * kv) * 128 // tp, hidden, 0, 1), (hidden, heads * 128 // tp, 0, 0), (inter // tp, hidden, 1, 1),
                       (hidden, inter // tp, 0, 0), (vocab // tp, hidden, 0, 1)]
    out = (ctypes.c_int * 12)()
    for (N, K, pair, norm) in shapes:
        if K % 64:
            continue
        for M in (1, 8):
            rc = lib.tdist_mega_linear_plan(M, N, K, pair, norm, 148, ctypes.byref(out))
            assert rc == 0, (M, N, K)
            cpt, tiles, nblk, R, nsub, kstage, kboxes, stages, scratch, bufs, slot, limit = list(out)
            assert tiles * cpt >= N and (tiles - 1) * cpt < N, (N, K, cpt, tiles)
            assert 1 <= R <= 24 and nblk * R >= cpt and (nblk - 1) * R < cpt
            assert R * nsub * 128 <= slot and kstage == 64 * nsub and kboxes * kstage >= K and (kboxes - 1) * kstage < K
            assert stages == (2 if pair else 1) * nblk * kboxes
            assert cpt <= 256 and tiles <= 148 * ((N + 148 * 256 - 1) // (148 * 256))
            if M == 1:
                assert bufs >= 8, (N, K, scratch, bufs)  # decode batch 1 always keeps at least 8 stage buffers
            # the staged activations [M][K] bf16 share the SM with the stage buffers: batch 8 x K 8192 is over the limit
            # and compile() refuses it with a clear message; everything up to ~100 KB of activations must fit
            assert bufs >= 2 or M * K * 2 > 100000, (M, N, K, scratch, bufs)


def test_contextual_autotune_picks_fastest_config():
    """Two tunables inside one thunk: the sweep walks the product of their spaces and caches the fastest combination."""
    import time

    from triton_dist.autotuner import contextual_autotune, tunable

    calls = []

    @tunable("t_a", {"x": [3, 1, 2]}, key=lambda *a, **k: "k")
    def fa(x=0):
        calls.append(("a", x))
        time.sleep(0.02 * x)  # (steps far above the scheduling noise of a loaded box)

    @tunable("t_b", {"y": [2, 0]}, key=lambda *a, **k: "k")
    def fb(y=5):
        calls.append(("b", y))
        time.sleep(0.02 * y)

    def thunk():
        fa()
        fb()

    contextual_autotune(is_dist=False, n_warmup=0, n_repeat=2)(thunk)()
    assert fa.best["k"] == {"x": 1} and fb.best["k"] == {"y": 0}
    assert {c for c in calls if c[0] == "a"} == {("a", 1), ("a", 2), ("a", 3)}
    calls.clear()
    thunk()  # outside a session the cached winners are used
    assert calls == [("a", 1), ("b", 0)]


def test_kernel_tile_walk_matches_python_restatement():
    """The tile_coord the GEMM kernel calls (exposed on the host) covers every tile exactly once, visits the shards in the
    documented order, and agrees with the Python restatements in kernels/nvidia/*_threadblock_swizzle.py."""
    import ctypes

    from triton_dist import _lib
    from triton_dist.kernels.nvidia.ag_gemm_threadblock_swizzle import ag_gemm_tile_order
    from triton_dist.kernels.nvidia.gemm_rs_threadblock_swizzle import gemm_rs_tile_order

    lib = _lib.load()
    if lib is None or not hasattr(lib, "tdist_gemm_tile_order"):
        import pytest

        pytest.skip("native library not built")
    for (M, N, W, bn, gm) in [(8192, 14336, 8, 256, 8), (8192, 4096, 2, 256, 8), (1600, 1000, 4, 192, 3), (4096, 1792, 1, 224, 8)]:
        for rank in range(W):
            for mode, ref in ((1, ag_gemm_tile_order), (2, gemm_rs_tile_order)):
                cap = 1 << 16
                out = (ctypes.c_int * (3 * cap))()
                n = lib.tdist_gemm_tile_order(mode, M, N, W, rank, bn, gm, out, cap)
                got = [(out[3 * t], out[3 * t + 1], out[3 * t + 2]) for t in range(n)]
                assert len(set(got)) == n == len(got)
                assert got == ref(M, N, W, rank, block_m=128, block_n=bn, group_m=gm), (M, N, W, rank, mode)
                first = (rank + (1 if mode == 2 else 0)) % W
                assert got[0][0] == first and got[-1][0] == (rank if mode == 2 else (rank + W - 1) % W)
    # plain mode: a permutation of the tile grid
    out = (ctypes.c_int * (3 * 4096))()
    n = lib.tdist_gemm_tile_order(0, 1000, 1032, 1, 0, 128, 8, out, 4096)
    got = {(out[3 * t + 1], out[3 * t + 2]) for t in range(n)}
    assert n == 8 * 9 and got == {(m, c) for m in range(8) for c in range(9)}


def test_launch_metadata_labels(monkeypatch):
    """TDIST_NVTX=1 wraps an op in an NVTX range named after the op with its shape / flops / bytes fields (counterpart of the
    reference kernels' launch_metadata); without the variable the decorator returns the function itself."""
    import torch

    import triton_dist.utils as U

    def op(a, b):
        return a.shape[0] + b.shape[0]

    assert U.launch_metadata("gemm")(op) is op or U._NVTX  # default: no wrapper at all
    labels = []
    monkeypatch.setattr(U, "_NVTX", True)
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda.nvtx, "range_push", lambda s: labels.append(("push", s)))
    monkeypatch.setattr(torch.cuda.nvtx, "range_pop", lambda: labels.append(("pop", None)))
    wrapped = U.launch_metadata("gemm", lambda a, b: U.gemm_metadata(a.shape[0], b.shape[0], a.shape[1]))(op)
    assert wrapped(torch.zeros(4, 8), torch.zeros(6, 8)) == 10
    assert labels == [("push", "gemm[M=4,N=6,K=8,W=1,flops=384,bytes=208,link_bytes=0]"), ("pop", None)]
    # a describe() that fails must not break the op
    labels.clear()
    assert U.launch_metadata("x", lambda *a: 1 / 0)(op)(torch.zeros(1, 1), torch.zeros(2, 1)) == 3
    assert labels[0] == ("push", "x")


class _FakeTensor:
    def __init__(self, ptr):
        self._ptr = ptr

    def data_ptr(self):
        return self._ptr

    def zero_(self):
        return self


def _bare_heap(monkeypatch, nbytes):
    """A SymmetricHeap object with only its allocator state (no CUDA): alloc / free are pure offset arithmetic."""
    import torch

    from triton_dist.runtime import symm

    heap = object.__new__(symm.SymmetricHeap)
    heap.rank, heap.world, heap.nbytes = 0, 1, nbytes
    heap.base = [1 << 40]
    heap.device = torch.device("cpu")
    heap._free, heap._live = [(0, nbytes)], {}
    monkeypatch.setattr(symm, "tensor_from_ptr", lambda ptr, shape, dtype, device, owner=None: _FakeTensor(ptr))
    return heap


def test_symmetric_heap_allocator_invariants(monkeypatch):
    """Random alloc / free sequences on the heap's first-fit allocator: live blocks never overlap, stay inside the heap and
    are 256-byte aligned; freeing everything merges the free list back into one block; exhaustion raises MemoryError.
    (Every rank runs the same sequence, so identical behaviour = identical offsets = symmetric addresses.)"""
    import random

    import torch

    nbytes = 1 << 20
    heap = _bare_heap(monkeypatch, nbytes)
    rng = random.Random(7)
    live = {}
    for step in range(3000):
        if live and (rng.random() < 0.45 or len(live) > 64):
            ptr = rng.choice(list(live))
            heap.free(_FakeTensor(ptr))
            del live[ptr]
        else:
            n = rng.choice([1, 7, 256, 1000, 4096, 40000])
            try:
                t = heap.alloc((n,), torch.uint8, zero=bool(rng.getrandbits(1)))
            except MemoryError:
                assert sum(live.values()) + n > nbytes // 2  # only a (fragmented) nearly full heap may refuse
                continue
            off = t.data_ptr() - heap.base[0]
            size = (n + 255) // 256 * 256
            assert off % 256 == 0 and 0 <= off and off + size <= nbytes
            for o, s in live.items():
                assert off + size <= o - heap.base[0] or o - heap.base[0] + s <= off, "live blocks overlap"
            live[t.data_ptr()] = size
        free_bytes = sum(s for _, s in heap._free)
        assert free_bytes + sum(live.values()) == nbytes
        assert all(a[0] + a[1] < b[0] for a, b in zip(heap._free, heap._free[1:])), "free list not sorted / not merged"
    for ptr in list(live):
        heap.free(_FakeTensor(ptr))
    assert heap._free == [(0, nbytes)] and not heap._live
    import pytest

    with pytest.raises(MemoryError):
        heap.alloc((nbytes + 1,), torch.uint8)
    heap.free(_FakeTensor(12345))  # unknown pointer: ignored
    assert heap._free == [(0, nbytes)]


def test_symmetric_heap_same_sequence_same_offsets(monkeypatch):
    """Two ranks that run the same alloc / free sequence get the same offsets: the property that makes ``symm_at`` pure
    pointer arithmetic."""
    import torch

    seqs = []
    for rank_base in (1 << 40, 5 << 40):
        heap = _bare_heap(monkeypatch, 1 << 20)
        heap.base = [rank_base]
        offs, held = [], []
        for i, n in enumerate([100, 5000, 300, 70000, 256, 9000]):
            t = heap.alloc((n,), torch.uint8)
            held.append(t)
            offs.append(t.data_ptr() - rank_base)
            if i % 2 == 1:
                heap.free(held.pop(0))
        seqs.append(offs)
    assert seqs[0] == seqs[1]


def test_perf_models_reproduce_the_bench_roofline():
    """gemm_perf_model / comm_perf_model give the roofline that bench.py and profiles/README.md quote for BASELINE config 2
    at TP8: 0.0763 ms per op = 56 MiB over 770 GB/s (the transfer bounds it, not the FLOPs)."""
    import torch

    from triton_dist.kernels.nvidia.comm_perf_model import (estimate_all_gather_time_ms, estimate_all_reduce_time_ms,
                                                            estimate_reduce_scatter_time_ms, fused_op_roofline_ms, get_nvlink_gbps)
    from triton_dist.kernels.nvidia.gemm_perf_model import estimate_gemm_sol_time_ms, get_dram_gbps, get_tensorcore_tflops

    M, K, N, W = 8192, 4096, 14336, 8
    link_bytes = (W - 1) * (M // W) * K * 2
    t_ag = estimate_all_gather_time_ms((M // W) * K * 2, W, gbps=770.0)
    assert abs(t_ag - link_bytes / (770.0 * 1e6)) < 1e-9 and abs(t_ag - 0.0763) < 5e-4
    assert abs(estimate_reduce_scatter_time_ms(M * 4096 * 2, W, gbps=770.0) - 0.0763) < 5e-4  # the down projection's partial
    flops_per_gpu = 2.0 * M * (N // W) * K
    assert abs(fused_op_roofline_ms(flops_per_gpu, link_bytes, 1699.5, 770.0) - 0.0763) < 5e-4  # link-bound
    assert fused_op_roofline_ms(flops_per_gpu, 0, 1699.5) == flops_per_gpu / (1699.5 * 1e9)      # compute-bound without traffic
    assert estimate_all_reduce_time_ms(1 << 26, W, 770.0, nvls=True) < estimate_all_reduce_time_ms(1 << 26, W, 770.0, nvls=False)
    assert get_nvlink_gbps(measured=False) == 900.0
    # nominal tables: B200 dense bf16 2.25 PFLOP/s, fp8 twice that; a large square GEMM is compute-bound, a GEMV memory-bound
    assert get_tensorcore_tflops(torch.bfloat16, "NVIDIA B200", measured=False) == 2250.0
    assert get_tensorcore_tflops(torch.float8_e4m3fn, "NVIDIA B200", measured=False) == 4500.0
    assert get_dram_gbps("NVIDIA B200") == 7700.0
    big = estimate_gemm_sol_time_ms(8192, 8192, 8192, torch.bfloat16, "NVIDIA B200")
    assert abs(big - 2.0 * 8192**3 / (2250.0 * 1e9)) < 1e-9
    gemv = estimate_gemm_sol_time_ms(1, 5120, 25600, torch.bfloat16, "NVIDIA B200")
    assert abs(gemv - (25600 + 5120 * 25600 + 5120) * 2 / (7700.0 * 1e6)) < 1e-9
