"""`python -m triton_dist.tools.compile_aot [--list] [--build]`

Counterpart of the reference's `tools/compile_aot.py` (`aot_compile_spaces` tags a Triton kernel with the signatures /
grids / algo-infos to pre-compile, `scripts/gen_aot_code.sh` emits C launchers). Here nothing is JIT-compiled:

* `build_library()` = `build.py` (nvcc, sm_100a) -> `triton_dist/lib/libtdist.so`;
* `list_kernels()` returns the template-instantiation table that is inside the library (kernel name -> instantiations),
  read from the symbol table — the native analogue of the reference's `__triton_algo_info_t` lists;
* `aot_compile_spaces` keeps the decorator name so that code written against the reference imports; it only records its
  argument on the function.
"""
from __future__ import annotations

import argparse
import re
import subprocess
from collections import defaultdict
from pathlib import Path
from typing import Dict, List

from .. import _lib


def aot_compile_spaces(spaces: dict):
    assert isinstance(spaces, dict)

    def deco(fn):
        fn.__aot_compile_spaces__ = spaces
        return fn

    return deco


def build_library(verbose: bool = False, clean: bool = False) -> Path:
    import importlib.util

    root = Path(__file__).resolve().parents[2]
    spec = importlib.util.spec_from_file_location("tdist_build", root / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return Path(mod.build(verbose, clean))


def list_kernels() -> Dict[str, List[str]]:
    """Kernel family -> list of instantiations (demangled), from `cuobjdump -res-usage` of the built library."""
    lib = _lib.lib_path()
    if not lib.exists():
        raise FileNotFoundError(f"{lib} is not built (python build.py)")
    out = subprocess.run(["cuobjdump", "-res-usage", str(lib)], capture_output=True, text=True, check=True).stdout
    names = re.findall(r"Function (\S+):", out)
    if names:
        dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout.split("\n")
    else:
        dem = []
    table: Dict[str, List[str]] = defaultdict(list)
    for d in dem:
        d = d.strip()
        if not d:
            continue
        d = re.sub(r"^void ", "", d)
        fam = re.sub(r"<.*", "", d.split("(")[0])
        table[fam].append(d.split("(")[0])
    return dict(table)


def exported_entry_points() -> List[str]:
    """C ABI entry points (`tdist_*`) of the library."""
    out = subprocess.run(["nm", "-D", "--defined-only", str(_lib.lib_path())], capture_output=True, text=True, check=True).stdout
    return sorted(line.split()[-1] for line in out.splitlines() if " T tdist_" in line)


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--build", action="store_true")
    ap.add_argument("--list", action="store_true")
    a = ap.parse_args()
    if a.build:
        print(build_library())
    if a.list or not a.build:
        for fam, inst in sorted(list_kernels().items()):
            print(f"{fam}: {len(inst)} instantiation(s)")
            for i in inst:
                if i != fam:
                    print(f"    {i}")
        print(f"{len(exported_entry_points())} C entry points (include/tdist.h)")
