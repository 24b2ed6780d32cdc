"""Same names as the reference's ``kernels/nvidia/common_ops.py`` (barriers, signal set / wait, bisect helpers); the
implementation lives in ``triton_dist.kernels.common_ops`` (host launchers) and ``csrc/device/symm.cuh`` (device)."""
from ..common_ops import *  # noqa: F401,F403
from ..common_ops import BarrierAllContext, CollectiveState, barrier_all_on_stream, set_signal, wait_eq  # noqa: F401
