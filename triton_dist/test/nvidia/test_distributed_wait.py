"""Counterpart of the reference's test/nvidia/test_distributed_wait.py: runs the notify + barrier check(s) of test_primitives.
    torchrun --nproc-per-node N -m triton_dist.test.nvidia.test_distributed_wait"""
from triton_dist.test.nvidia.test_primitives import run

if __name__ == "__main__":
    run(['notify', 'barrier'])
