#!/bin/bash
# compute-sanitizer passes over the single-GPU kernel pieces (the reference only documents
# `compute-sanitizer --tool memcheck torchrun ...` in a comment of its launch.sh:165-167).
#   bash scripts/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest -k expression]
# Run ONE tool per invocation on ONE GPU; kernels that spin on peers are excluded (the sanitizer serialises CTAs, which
# deadlocks producer/consumer protocols), so this covers GEMM (plain / grouped), decode kernels, MoE helpers, SP attention.
set -e
TOOL=${1:-memcheck}
EXPR=${2:-"gemm_matches or tile_width or skinny or decode or rmsnorm or gemv or moe_align or grouped or noalloc"}
export PYTHONPATH=$PWD:$PYTHONPATH
exec compute-sanitizer --tool "$TOOL" --error-exitcode 1 --launch-timeout 0 --target-processes all \
  python -m pytest tests/test_gpu_ops.py tests/test_gpu_decode.py tests/test_gpu_moe.py tests/test_zz_no_alloc.py -x -q -m gpu -k "$EXPR"
