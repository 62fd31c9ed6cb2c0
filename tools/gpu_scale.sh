#!/usr/bin/env bash
# scaling point at N GPUs: our bench (fused TP), optionally nccl + reference
N=${1:-8}; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811"
for impl in "$@"; do
  case $impl in
    fused|nccl) timeout 300 $TR bench.py --gpus $N --steps 64 --warmup 8 --tp-impl $impl > gpurun_out/scale_${impl}_$N.log 2>&1; echo "$impl rc=$?";;
    ref) timeout 300 $TR bench.py --gpus $N --steps 64 --warmup 8 --impl reference > gpurun_out/scale_ref_$N.log 2>&1; echo "ref rc=$?";;
    tests) timeout 300 python -m pytest tests/test_parallel_gpu.py -x -q --timeout 200 -p no:cacheprovider > gpurun_out/scale_tests_$N.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/scale_tests_$N.log;;
  esac
  grep -h '^{' gpurun_out/scale_${impl}_$N.log 2>/dev/null | cut -c1-700
  grep -hE "Error|error:" gpurun_out/scale_${impl}_$N.log 2>/dev/null | head -5
done
