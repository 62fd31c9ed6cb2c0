#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -x -q 2>&1 | tail -3
for Q in Q8_0; do python bench.py --ftype $Q --steps 128 --warmup 16 2>/dev/null | grep '^{' > gpurun_out/bench_$Q.json; python -c "import json; d=json.load(open('gpurun_out/bench_$Q.json')); print('$Q tg', d['value'], 'tok/s', d['ms_per_step'], 'ms  prefill', d['prefill_tok_s'], 'e2e', d['e2e']['value'])"; done
