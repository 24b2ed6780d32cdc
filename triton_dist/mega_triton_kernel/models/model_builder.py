"""Same import path as the reference's ``mega_triton_kernel/models/model_builder.py``."""
from ..core.builder import NUM_KV_SPLITS_DEFAULT, ModelBuilder  # noqa: F401
