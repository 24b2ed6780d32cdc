#!/bin/bash
# Final N-GPU (default 8) run: headline bench first, then correctness + perf of the fused ops, megakernel, collectives, EP.
# Every step has its own hard timeout; if the CTA-pair GEMM misbehaves at this world size the rest runs on single-CTA tiles.
OUT=${1:-gpurun_out}
N=${2:-8}
mkdir -p $OUT
export PYTHONPATH=$PWD:$PYTHONPATH
step() { name=$1; tmo=$2; shift 2; timeout -k 5 $tmo "$@" > $OUT/$name.log 2>&1; rc=$?; echo "== $name rc=$rc"; grep -E "passed|failed|^PASS|^FAIL|suite\]|speedup|metric|Error|ok$|FAILED|ms/token|busbw|dispatch " $OUT/$name.log | grep -v Traceback | cut -c1-${CUT:-400} | tail -${TAILN:-6}; return $rc; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
step ag_w$N 50 $TR --master-port 29741 -m triton_dist.test.nvidia.test_ag_gemm 8192 14336 4096 --verify-iters 2 || export TDIST_GEMM_CTA_GROUP=1
step rs_w$N 50 $TR --master-port 29742 -m triton_dist.test.nvidia.test_gemm_rs 8192 4096 14336 --verify-iters 2 || export TDIST_GEMM_CTA_GROUP=1
echo "TDIST_GEMM_CTA_GROUP=${TDIST_GEMM_CTA_GROUP:-default}"
CUT=4000 step bench_w$N 150 $TR --master-port 29743 bench.py --gpus $N --steps 20 --warmup 5
step mega_w$N 80 $TR --master-port 29744 -m triton_dist.mega_triton_kernel.test.models.test_qwen3
step ep_w$N 60 $TR --master-port 29746 -m triton_dist.test.nvidia.test_ep_a2a -M 128 -N 7168 -G 256 --topk 8 --fp8 --verify-iters 1
step gemm_ar_w$N 60 $TR --master-port 29747 -m triton_dist.test.nvidia.test_gemm_ar 2048 5120 25600 --verify-iters 1
TAILN=40 step coll_w$N 90 $TR --master-port 29745 -m triton_dist.benchmark.bench_collectives --max-bytes 67108864 --iters 10 --json $OUT/collectives_w$N.json
