#!/bin/bash
# Launch a distributed script or module on all GPUs of this node (reference scripts/launch.sh:128-179).
#   bash scripts/launch.sh [--nproc_per_node=N] <script.py | -m package.module> [args...]
# One process per GPU over torch.distributed (NCCL for glue, the symmetric heap for everything else).
set -e
SCRIPT_DIR=$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)
export PYTHONPATH="${SCRIPT_DIR}/..:${PYTHONPATH}"
export CUDA_DEVICE_MAX_CONNECTIONS=${CUDA_DEVICE_MAX_CONNECTIONS:-1}
export TDIST_SYMMETRIC_SIZE=${TDIST_SYMMETRIC_SIZE:-${NVSHMEM_SYMMETRIC_SIZE:-4294967296}}
export NCCL_DEBUG=${NCCL_DEBUG:-ERROR}
nproc_per_node=$(nvidia-smi -L 2>/dev/null | wc -l)
[ "$nproc_per_node" -ge 1 ] || nproc_per_node=1
if [[ "$1" == --nproc_per_node=* ]]; then
  nproc_per_node="${1#--nproc_per_node=}"
  shift
fi
MASTER_PORT=${MASTER_PORT:-$((23000 + RANDOM % 2000))}
# for race hunting: compute-sanitizer --tool memcheck python -m torch.distributed.run ... (single-GPU pieces only)
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node="${nproc_per_node}" \
  --master-addr 127.0.0.1 --master-port "${MASTER_PORT}" ${DIST_TRITON_EXTRA_TORCHRUN_ARGS} "$@"
