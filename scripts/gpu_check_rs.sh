#!/bin/bash
# GEMM+RS focused check on N GPUs (default 2)
OUT=${1:-gpurun_out}
N=${2:-2}
mkdir -p $OUT
export PYTHONPATH=$PWD:$PYTHONPATH
step() { name=$1; tmo=$2; shift 2; timeout $tmo "$@" > $OUT/$name.log 2>&1; rc=$?; echo "== $name rc=$rc"; grep -E "passed|failed|PASS|FAIL|suite\]|speedup|metric|Error|err |ok$|FAILED" $OUT/$name.log | grep -v Traceback | cut -c1-${CUT:-300} | tail -${TAILN:-10}; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
step rs_small 120 $TR --master-port 29721 -m triton_dist.test.nvidia.test_gemm_rs $((2048*N)) 1024 $((512*N)) --check --stress
step rs_odd 120 $TR --master-port 29722 -m triton_dist.test.nvidia.test_gemm_rs $((200*N)) 136 $((264*N)) --check --verify-iters 2
step rs_pair 120 $TR --master-port 29717 -m triton_dist.test.nvidia.test_gemm_rs 8192 4096 14336 --verify-iters 2 --stress
TDIST_GEMM_CTA_GROUP=1 step rs_single 120 $TR --master-port 29719 -m triton_dist.test.nvidia.test_gemm_rs 8192 4096 14336 --verify-iters 2
CUT=1500 step bench 200 $TR --master-port 29720 bench.py --gpus $N --steps 10 --warmup 3 --no-decode
