"""Counterpart of the reference's test/nvidia/test_notify.py: runs the notify check(s) of test_primitives.
    torchrun --nproc-per-node N -m triton_dist.test.nvidia.test_notify"""
from triton_dist.test.nvidia.test_primitives import run

if __name__ == "__main__":
    run(['notify'])
