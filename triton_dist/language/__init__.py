"""Device-language surface of the reference, documented for the native framework.

In the reference these are Triton builtins lowered by an MLIR dialect (``language/distributed_ops.py:56-111``,
``language/simt_ops.py:29-67``, ``language/extra/libshmem_device.py``). Here kernels are CUDA C++, so the primitives are
``__device__`` functions in ``csrc/device/symm.cuh`` / ``csrc/device/ptx.cuh``. This module maps every reference name
to its C++ counterpart (``PRIMITIVES``) so that code reading the reference can find the native equivalent, and exposes
the host-side constants that share values with NVSHMEM."""
from .core import PRIMITIVES, describe  # noqa: F401
from .extra import libshmem_device  # noqa: F401

NVSHMEM_SIGNAL_SET = 9
NVSHMEM_SIGNAL_ADD = 10
