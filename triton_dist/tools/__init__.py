"""Ahead-of-time tooling. The reference needs an AOT pipeline (tools/compile_aot.py, tools/compile, tools/runtime) to turn
JIT Triton kernels into C launchers; every kernel of this build is compiled ahead of time by nvcc into libtdist.so, so
what remains is the instantiation table and the build entry point."""
from .compile_aot import aot_compile_spaces, build_library, list_kernels  # noqa: F401
