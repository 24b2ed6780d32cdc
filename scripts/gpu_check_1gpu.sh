#!/bin/bash
# Single-GPU validation with a hard wall-clock limit per step (a hung kernel cannot eat the GPU budget).
OUT=${1:-gpurun_out}
mkdir -p $OUT
export PYTHONPATH=$PWD:$PYTHONPATH
step() { name=$1; tmo=$2; shift 2; timeout $tmo "$@" > $OUT/$name.log 2>&1; rc=$?; echo "== $name rc=$rc"; grep -E "passed|failed|error|PASS|FAIL|suite\]|smoke|metric|Error" $OUT/$name.log | grep -v Traceback | tail -${TAILN:-8}; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=1 --master-addr 127.0.0.1"
step decode 240 python -m pytest tests/test_gpu_decode.py -x -q
step noalloc 90 python -m pytest tests/test_zz_no_alloc.py -x -q
step moe_kernels 150 python -m pytest tests/test_gpu_moe.py -x -q -k "align or grouped"
TAILN=14 step suite_moe 240 $TR --master-port 29701 -m triton_dist.test.nvidia.run_suite --suite moe
TAILN=14 step suite_models 300 $TR --master-port 29702 -m triton_dist.test.nvidia.run_suite --suite models
step smoke 120 python -c "import __graft_entry__ as g; g.smoke()"
step bench1 150 python bench.py --steps 5 --warmup 3
