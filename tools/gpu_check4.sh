#!/usr/bin/env bash
# full GPU suite + Q4_K_M bench (mixed-type QKV kept in block types)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --ftype Q4_K_M --steps 64 --warmup 8 2>/dev/null | grep '^{' > gpurun_out/bench_q4km.json
python -c "import json; d=json.load(open('gpurun_out/bench_q4km.json')); print('Q4_K_M tg', d['value'], 'tok/s  prefill', d['prefill_tok_s'], 'e2e', d['e2e']['value'])"
