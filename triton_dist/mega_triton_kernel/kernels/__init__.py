"""Task bodies. The reference keeps one Triton device function per task type in kernels/{linear,mlp_fc1,norm,activation,
elementwise,flash_attn,allreduce,barrier,prefetch}.py; the native bodies are the ``*_body`` functions of
csrc/mega/megakernel.cu (see ``core.registry`` for the mapping). ``torch_impl`` holds the eager equivalents the builder
runs on a machine without a GPU and the tests use as oracles."""
from ..test import torch_impl_utils as torch_impl  # noqa: F401
