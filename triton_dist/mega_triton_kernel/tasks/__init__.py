"""Task definitions. The reference has one Python class per task type (tasks/{linear,norm,activation,elementwise,attn,
allreduce,barrier,prefetch}.py) that computes tiling and dependencies; natively that is ``tiles_of`` / ``smem_of`` /
``tdist_mega_add_op`` in csrc/mega/builder.cc. The op table is exposed as ``core.registry.REGISTRY``."""
from ..core.registry import BY_CODE, REGISTRY, OpInfo, op_name  # noqa: F401
from ..core.task_base import Task  # noqa: F401
