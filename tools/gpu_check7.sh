#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -k "gemv" -x -q 2>&1 | tail -5
timeout 600 python -m pytest tests/test_engine_gpu.py -x -q 2>&1 | tail -3
for CFG in "12 0" "0 0" "25 0" "12 1"; do set -- $CFG
  DLP_GEMV_POOL=$1 DLP_ATTN_ONEPASS=$2 python bench.py --steps 64 --warmup 8 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('pool=$1 onepass=$2  F16 tg', d['value'], 'tok/s', d['ms_per_step'], 'ms')"
done
DLP_GEMV_POOL=12 timeout 300 python tools/timeline_cta.py --layers 4 2>&1 | tail -12
