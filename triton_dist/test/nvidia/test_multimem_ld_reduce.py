"""Counterpart of the reference's test/nvidia/test_multimem_ld_reduce.py: runs the multimem check(s) of test_primitives.
    torchrun --nproc-per-node N -m triton_dist.test.nvidia.test_multimem_ld_reduce"""
from triton_dist.test.nvidia.test_primitives import run

if __name__ == "__main__":
    run(['multimem'])
