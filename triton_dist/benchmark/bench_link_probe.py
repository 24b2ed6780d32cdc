"""Bandwidth probes of the NVLink / NVLS data paths the fused kernels are built from (csrc/kernels/link_probe.cu).
    torchrun --nproc-per-node N -m triton_dist.benchmark.bench_link_probe [--mb 8] [--json out.json] [--bulk]
Per mode, grid and loads in flight: device time (max over ranks) and GB/s per GPU of the bytes that cross the link in the
limiting direction. No counterpart in the reference (its comm_perf_model.py reads link speeds from tables)."""
import argparse
import ctypes
import json

import torch
import torch.distributed as dist

from triton_dist import _lib
from triton_dist.runtime import get_heap
from triton_dist.test.common import finish, setup
from triton_dist.utils import dist_print, symm_tensor

MODES = {0: "local copy", 1: "nvls multimem.st push", 2: "unicast push x(W-1)", 3: "pull from W-1 peers",
         4: "nvls multimem.ld_reduce", 5: "bulk ring -> multicast", 6: "bulk ring pull"}


def link_bytes(mode: int, nbytes: int, world: int) -> int:
    """bytes through the busier link direction of one GPU"""
    if mode == 0:
        return nbytes
    if mode in (1, 5):
        return nbytes * world  # ingress: every rank receives all W shards (its own one comes back from the switch)
    if mode == 4:
        return nbytes * world  # egress: every rank serves its copy of all W segments
    return nbytes * (world - 1)


def run(mode, src, symm, out, nbytes, grid, threads, unroll, iters):
    lib = _lib.lib()
    lib.tdist_link_probe.restype = ctypes.c_int
    lib.tdist_link_probe.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    heap = get_heap()

    def call():
        _lib.check(lib.tdist_link_probe(ctypes.byref(heap.ctx), mode, src.data_ptr(), symm.data_ptr(), out.data_ptr(), nbytes,
                                        grid, threads, unroll, _lib.stream_ptr()), "link_probe")

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    heap.barrier_all()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        call()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=8, help="MiB per rank")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--json", default="")
    ap.add_argument("--bulk", action="store_true", help="also run the bulk-copy rings (mode 5 stores to a multicast address with cp.async.bulk)")
    args = ap.parse_args()
    group, rank, world = setup()
    nbytes = args.mb << 20
    src = torch.randn(nbytes // 2, device="cuda").bfloat16()
    symm = symm_tensor((world, nbytes // 2), torch.bfloat16)
    symm.normal_()
    out = torch.empty(world * nbytes // 2, device="cuda", dtype=torch.bfloat16)
    get_heap().barrier_all()
    torch.cuda.synchronize()
    rows = []
    sweep = [(0, [(148, 512, 8)])]
    if world > 1:
        sweep += [(1, [(2, 256, 8), (4, 256, 8), (8, 256, 8), (16, 256, 8), (32, 256, 8), (32, 512, 16), (148, 512, 8)]),
                  (2, [(8, 256, 8), (16, 256, 8), (32, 256, 8), (32, 512, 16), (148, 512, 8)]),
                  (3, [(8, 256, 8), (16, 256, 16), (32, 256, 16), (148, 64, 8), (148, 64, 16), (148, 512, 8)]),
                  (4, [(4, 256, 8), (8, 256, 16), (16, 256, 16), (32, 256, 16), (148, 64, 8), (148, 64, 16), (148, 512, 8)])]
        if args.bulk:
            sweep += [(6, [(4, 32, 0), (8, 32, 0), (16, 32, 0), (32, 32, 0)]), (5, [(1, 32, 0), (2, 32, 0), (4, 32, 0), (8, 32, 0)])]
    for mode, cfgs in sweep:
        if mode in (1, 4, 5) and not get_heap().multicast:
            continue
        if mode == 5:  # wipe what mode 1 delivered, so that the check below really tests the bulk store
            torch.cuda.synchronize()
            get_heap().barrier_all()
            symm.zero_()
            torch.cuda.synchronize()
            get_heap().barrier_all()
        for grid, threads, unroll in cfgs:
            ms = run(mode, src, symm, out, nbytes, grid, threads, unroll, args.iters)
            gbs = link_bytes(mode, nbytes, world) / ms / 1e6
            rows.append(dict(mode=mode, name=MODES[mode], grid=grid, threads=threads, unroll=unroll, us=ms * 1e3, gbs=gbs))
            dist_print(f"mode {mode} {MODES[mode]:28s} grid {grid:4d} x {threads:3d} thr, {unroll:2d} in flight: {ms * 1e3:8.1f} us  {gbs:7.1f} GB/s")
    ok = True
    if args.bulk and world > 1 and get_heap().multicast:
        # did the bulk store to the multicast address deliver? every rank's slot must hold that rank's src
        get_heap().barrier_all()
        torch.cuda.synchronize()
        mine = [torch.empty_like(src) for _ in range(world)]
        dist.all_gather(mine, src)
        ok = all(bool(torch.equal(symm[r], mine[r])) for r in range(world))
        dist_print(f"bulk store to multicast address delivered everywhere: {ok}", allowed_ranks="all")
    if args.json and rank == 0:
        with open(args.json, "w") as f:
            json.dump(dict(world=world, mb=args.mb, rows=rows), f, indent=1)
    finish(ok)
