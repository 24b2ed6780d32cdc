"""Layer-level graph helpers (reference models/layers): the Qwen3 decoder layer as a sequence of ModelBuilder calls."""
from ..qwen3 import Qwen3Model  # noqa: F401
