#!/usr/bin/env bash
# full GPU suite + attention-variant sweep + default bench (one gpurun call)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/ab_onepass.py 2>&1 | grep ctx
python bench.py --steps 128 --warmup 16 2>/dev/null | grep '^{' > gpurun_out/bench_f16.json
python -c "import json; d=json.load(open('gpurun_out/bench_f16.json')); print('F16 tg', d['value'], 'tok/s', d['ms_per_step'], 'ms prefill', d['prefill_tok_s'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])"
