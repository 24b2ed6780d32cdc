"""Counterpart of the reference's test/nvidia/test_common_ops.py: runs the barrier + ring_put check(s) of test_primitives.
    torchrun --nproc-per-node N -m triton_dist.test.nvidia.test_common_ops"""
from triton_dist.test.nvidia.test_primitives import run

if __name__ == "__main__":
    run(['barrier', 'ring_put'])
