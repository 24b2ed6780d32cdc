#!/usr/bin/env bash
# Populate baseline/_ref with the UNMODIFIED reference for `bench.py --impl reference`.
# 1. the documented offline install (fails: the reference has no root setup.py, and python/setup.py downloads LLVM and
#    an NVIDIA toolchain and builds a Triton fork -- see DESIGN.md section 6);
# 2. fallback: copy the reference's pure-Python package tree as it is. Its Triton kernels stay unbuildable, but its
#    torch/NCCL code paths (TP_MLP.torch_ag_gemm / torch_gemm_rs / torch_fwd) run unmodified.
set -u
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
DST="$REPO/baseline/_ref"
mkdir -p "$DST"
if python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$DST" "$REF" >"$DST/pip_install.log" 2>&1; then
  echo "pip install of the reference succeeded"; exit 0
fi
echo "pip install failed (log: baseline/_ref/pip_install.log); copying the unmodified python package tree"
rm -rf "$DST/triton_dist"
cp -r "$REF/python/triton_dist" "$DST/triton_dist"
(cd "$REF" && git rev-parse HEAD 2>/dev/null || echo unknown) > "$DST/REFERENCE_COMMIT"
echo "copied $(find "$DST/triton_dist" -name '*.py' | wc -l) files"
