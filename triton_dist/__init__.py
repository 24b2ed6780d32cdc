"""triton_dist (B200-native): compute/communication-overlapping kernels for
8xB200 (sm_100a, NVLink 5 / NVSwitch) with the Python API of Triton-distributed.

No Triton compiler and no NVSHMEM: hand-written CUDA kernels (tcgen05 / TMEM /
TMA / multimem) in ``libtdist.so`` over a VMM symmetric heap, PyTorch on top.
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401
