from .symm import SymmetricHeap, HostHeap, get_heap, init_heap, finalize_heap, heap_initialized, tensor_from_ptr  # noqa: F401
from .team import Team, team_world  # noqa: F401
