"""Runs the device-side self-test of the NVSHMEM-style primitives (csrc/kernels/primitives_selftest.cu): ring
putmem_signal + signal_wait_until, fcollect, broadcast (NVLS when bound), int_p, signal_op, fence / quiet, remote_ptr loads,
team_translate_pe — the device vocabulary of the reference's libshmem_device, called from a plain CUDA kernel.
    torchrun --nproc-per-node N -m triton_dist.test.nvidia.test_device_primitives"""
import ctypes

import torch

from triton_dist import _lib
from triton_dist.runtime import get_heap
from triton_dist.test.common import all_ok, base_parser, finish, setup, skew
from triton_dist.utils import dist_print, nvshmem_barrier_all_on_stream, symm_tensor

if __name__ == "__main__":
    p = base_parser(__doc__)
    args = p.parse_args()
    group, rank, world = setup()
    if not torch.cuda.is_available():
        dist_print("device primitives need a GPU: skipped")
        finish(True)
    heap = get_heap()
    buf = symm_tensor(((3 * world + 2) * 4,), torch.int32)
    flags = symm_tensor((2 * world + 1,), torch.int32)
    nvshmem_barrier_all_on_stream()
    ok = True
    for epoch in range(1, args.verify_iters + 1):
        skew(args.stress)
        _lib.check(_lib.lib().tdist_primitives_selftest(ctypes.byref(heap.ctx), buf.data_ptr(), flags.data_ptr(), epoch,
                                                        _lib.stream_ptr()), "primitives_selftest")
        torch.cuda.synchronize()
        got = int(buf[3 * world * 4].item())
        if got != epoch:
            dist_print(f"epoch {epoch}: self-test result {got}", allowed_ranks="all", prefix=True)
        ok = ok and got == epoch
        nvshmem_barrier_all_on_stream()
    finish(all_ok(ok))
