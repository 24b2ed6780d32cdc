"""The reference's test_simt.py exercises its SIMT dialect (`simt_exec_region`, `extract`, `insert`): scalar per-thread
code inside a tile program. CUDA kernels are SIMT to begin with, so there is nothing to emulate; the names are kept in
`triton_dist.language` with that note, and this script runs the primitive-map check."""
import runpy

if __name__ == "__main__":
    from triton_dist.language.core import describe

    for n in ("simt_exec_region", "extract", "insert"):
        print(describe(n))
    runpy.run_module("triton_dist.test.nvidia.test_language_extra", run_name="__main__")
