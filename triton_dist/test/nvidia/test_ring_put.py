"""Counterpart of the reference's test/nvidia/test_ring_put.py: runs the ring_put check(s) of test_primitives.
    torchrun --nproc-per-node N -m triton_dist.test.nvidia.test_ring_put"""
from triton_dist.test.nvidia.test_primitives import run

if __name__ == "__main__":
    run(['ring_put'])
