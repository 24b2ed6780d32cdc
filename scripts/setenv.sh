#!/bin/bash
# Environment for running scripts of this tree directly (counterpart of the reference's scripts/setenv.sh):
#   source scripts/setenv.sh
# scripts/launch.sh sets the same variables itself; this file is for `python some_script.py` / `torchrun ...` by hand.
_TDIST_ROOT=$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)
export PYTHONPATH="${_TDIST_ROOT}:${PYTHONPATH}"
# symmetric heap per rank (bytes); NVSHMEM_SYMMETRIC_SIZE is honoured as an alias
export TDIST_SYMMETRIC_SIZE=${TDIST_SYMMETRIC_SIZE:-${NVSHMEM_SYMMETRIC_SIZE:-4294967296}}
# one hardware queue per device keeps the launch order of the copy-engine producers and their consumers (reference launch.sh:128)
export CUDA_DEVICE_MAX_CONNECTIONS=${CUDA_DEVICE_MAX_CONNECTIONS:-1}
export NCCL_DEBUG=${NCCL_DEBUG:-ERROR}
export MASTER_ADDR=${MASTER_ADDR:-127.0.0.1}
if [ ! -f "${_TDIST_ROOT}/triton_dist/lib/libtdist.so" ]; then
  echo "libtdist.so is missing: run  python ${_TDIST_ROOT}/build.py" >&2
fi
unset _TDIST_ROOT
