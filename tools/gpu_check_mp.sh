#!/usr/bin/env bash
# multi-GPU check: parallel tests + bench at N GPUs for both TP implementations and the reference arm
N=${1:-2}
mkdir -p gpurun_out; : > gpurun_out/summary_mp.txt
run() { local name=$1 to=$2; shift 2; echo "=== $name ===" | tee -a gpurun_out/summary_mp.txt
  timeout "$to" "$@" > "gpurun_out/$name.log" 2>&1; local rc=$?
  echo "$name rc=$rc" | tee -a gpurun_out/summary_mp.txt; tail -n 15 "gpurun_out/$name.log"; }
run par_tests 700 python -m pytest tests/test_parallel_gpu.py -x -q --timeout 250 -p no:cacheprovider
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29801"
run bench_fused_$N 420 $TR bench.py --gpus $N --steps 64 --warmup 8 --tp-impl fused
run bench_nccl_$N 420 $TR bench.py --gpus $N --steps 64 --warmup 8 --tp-impl nccl
run bench_ref_$N 420 $TR bench.py --gpus $N --steps 64 --warmup 8 --impl reference
cat gpurun_out/summary_mp.txt
