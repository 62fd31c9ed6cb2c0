#!/usr/bin/env bash
mkdir -p gpurun_out
DLP_GEMV_PRE2=1 timeout 600 python -m pytest tests/test_kernels_gpu.py -k "gemv" tests/test_engine_gpu.py -x -q 2>&1 | tail -3
for PRE in 1 0 1 0; do
  DLP_GEMV_PRE2=$PRE python bench.py --steps 128 --warmup 16 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('pre2=$PRE  F16 tg', d['value'], 'tok/s', d['ms_per_step'], 'ms')"
done
