#!/bin/bash
# N-GPU validation (default 2) with hard wall-clock limits per step: a hung kernel cannot eat the GPU budget.
OUT=${1:-gpurun_out}
N=${2:-2}
mkdir -p $OUT
export PYTHONPATH=$PWD:$PYTHONPATH
step() { name=$1; tmo=$2; shift 2; timeout $tmo "$@" > $OUT/$name.log 2>&1; rc=$?; echo "== $name rc=$rc"; grep -E "passed|failed|PASS|FAIL|suite\]|speedup|metric|Error|err |TFLOP|ok$|FAILED|ms/token|ms_per" $OUT/$name.log | grep -v Traceback | cut -c1-${CUT:-300} | tail -${TAILN:-10}; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
TAILN=12 step suite_core 300 $TR --master-port 29711 -m triton_dist.test.nvidia.run_suite --suite core
TAILN=12 step suite_mega 300 $TR --master-port 29713 -m triton_dist.test.nvidia.run_suite --suite mega
TAILN=12 step suite_moe 300 $TR --master-port 29712 -m triton_dist.test.nvidia.run_suite --suite moe
TAILN=8 step suite_models 300 $TR --master-port 29714 -m triton_dist.test.nvidia.run_suite --suite models
step ag_pair 120 $TR --master-port 29716 -m triton_dist.test.nvidia.test_ag_gemm 8192 14336 4096 --verify-iters 2 --stress
step rs_pair 120 $TR --master-port 29717 -m triton_dist.test.nvidia.test_gemm_rs 8192 4096 14336 --verify-iters 2 --stress
export TDIST_GEMM_CTA_GROUP=1
step ag_single 120 $TR --master-port 29718 -m triton_dist.test.nvidia.test_ag_gemm 8192 14336 4096 --verify-iters 2
step rs_single 120 $TR --master-port 29719 -m triton_dist.test.nvidia.test_gemm_rs 8192 4096 14336 --verify-iters 2
unset TDIST_GEMM_CTA_GROUP
CUT=3000 step bench 300 $TR --master-port 29720 bench.py --gpus $N --steps 10 --warmup 3
