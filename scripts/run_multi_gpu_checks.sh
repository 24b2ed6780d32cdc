#!/bin/bash
# Multi-GPU validation + headline perf. usage: run_multi_gpu_checks.sh <ngpus> [outdir] [quick]
N=${1:-2}
OUT=${2:-gpurun_out}
QUICK=${3:-}
mkdir -p $OUT
export PYTHONPATH=$PWD:$PYTHONPATH
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
port=29600
run() { name=$1; shift; port=$((port+1)); timeout ${TMO:-150} $TR --master-port $port "$@" > $OUT/$name.log 2>&1; echo "$name rc=$?"; grep -E "PASS|FAIL|speedup|busbw|GB/s|setup|Error|error|metric" $OUT/$name.log | grep -v Traceback | tail -${TAILN:-12}; }
run ag_gemm_check_$N -m triton_dist.test.nvidia.test_ag_gemm $((2048*N)) $((512*N)) 1024 --check --stress
run gemm_rs_check_$N -m triton_dist.test.nvidia.test_gemm_rs $((2048*N)) 1024 $((512*N)) --check --stress
run gemm_ar_check_$N -m triton_dist.test.nvidia.test_gemm_ar 64 1024 $((512*N)) --check --low-latency --stress
run gemm_ar2_check_$N -m triton_dist.test.nvidia.test_gemm_ar 2048 1024 $((256*N)) --check
run ag_gemm_perf_$N -m triton_dist.test.nvidia.test_ag_gemm 8192 14336 4096
run gemm_rs_perf_$N -m triton_dist.test.nvidia.test_gemm_rs 8192 4096 14336
run bench_$N bench.py --gpus $N --steps 20 --warmup 5
if [ -z "$QUICK" ]; then
run gemm_ar_perf_$N -m triton_dist.test.nvidia.test_gemm_ar 128 5120 25600 --low-latency
TAILN=60 run allreduce_$N -m triton_dist.test.nvidia.test_allreduce --max-bytes $((256*1024*1024)) --verify-iters 1 --graph
TAILN=60 run fast_allgather_$N -m triton_dist.test.nvidia.test_fast_allgather --max-bytes $((4*1024*1024)) --verify-iters 1 --graph
run reduce_scatter_$N -m triton_dist.test.nvidia.test_reduce_scatter 8192 4096
fi
