#!/usr/bin/env python
"""Build libtdist.so (all CUDA kernels + native runtime) for sm_100a, in tree.

    python build.py            # incremental, parallel (ninja)
    python build.py --clean
    python build.py --verbose  # adds -Xptxas -v

The library is plain CUDA/C++ with a C ABI (include/tdist.h): no torch headers,
so a full rebuild takes well under a minute and the .so loads anywhere the CUDA
runtime does (cudart is linked statically, libcuda is resolved lazily).
"""
import argparse
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
SRC = ROOT / "csrc"
OUT_DIR = ROOT / "triton_dist" / "lib"
BUILD_DIR = ROOT / "build"
LIB = OUT_DIR / "libtdist.so"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-Xcompiler", "-Wno-unused-function", "--expt-relaxed-constexpr", f"-I{ROOT / 'include'}"]


def sources():
    cu = sorted((SRC / "kernels").glob("*.cu")) + sorted((SRC / "mega").glob("*.cu")) + sorted((SRC / "tutorials").glob("*.cu"))
    cc = sorted((SRC / "runtime").glob("*.cc")) + sorted((SRC / "mega").glob("*.cc"))
    return cu, cc


def write_ninja(verbose: bool) -> Path:
    cu, cc = sources()
    BUILD_DIR.mkdir(exist_ok=True)
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    flags = " ".join(ARCH + COMMON + (["-Xptxas", "-v"] if verbose else []))
    lines = [
        f"nvcc = {NVCC}",
        f"flags = {flags}",
        "rule cu",
        "  command = $nvcc $flags -MD -MF $out.d -c $in -o $out",
        "  depfile = $out.d",
        "  deps = gcc",
        "rule cc",
        "  command = $nvcc $flags -x cu -MD -MF $out.d -c $in -o $out",
        "  depfile = $out.d",
        "  deps = gcc",
        "rule link",
        "  command = $nvcc -shared -cudart static " + " ".join(ARCH) + " -o $out $in -lpthread -ldl -lrt",
    ]
    objs = []
    for s in cu:
        o = BUILD_DIR / (s.stem + ".cu.o")
        lines.append(f"build {o}: cu {s}")
        objs.append(str(o))
    for s in cc:
        o = BUILD_DIR / (s.stem + ".cc.o")
        lines.append(f"build {o}: cc {s}")
        objs.append(str(o))
    lines.append(f"build {LIB}: link {' '.join(objs)}")
    lines.append(f"default {LIB}")
    path = BUILD_DIR / "build.ninja"
    path.write_text("\n".join(lines) + "\n")
    return path


def build(verbose: bool = False, clean: bool = False) -> Path:
    if clean:
        shutil.rmtree(BUILD_DIR, ignore_errors=True)
        if LIB.exists():
            LIB.unlink()
    ninja_file = write_ninja(verbose)
    ninja = shutil.which("ninja")
    cmd = [ninja or sys.executable, "-f", str(ninja_file)] if ninja else [sys.executable, "-m", "ninja", "-f", str(ninja_file)]
    if verbose:
        cmd.append("-v")
    subprocess.run(cmd, check=True, cwd=str(BUILD_DIR))  # .ninja_log / .ninja_deps stay inside build/
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--clean", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--spin-limit", type=int, default=0, help="debug build: trap with a message after this many polls of one "
                    "flag (-DTDIST_SPIN_LIMIT); 0 = normal build")
    a = ap.parse_args()
    if a.spin_limit > 0:
        COMMON.append(f"-DTDIST_SPIN_LIMIT={a.spin_limit}")
        a.clean = True
    print(build(a.verbose, a.clean))
