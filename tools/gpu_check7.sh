#!/usr/bin/env bash
mkdir -p gpurun_out
DLP_SPLIT_V=1 python bench.py --ftype Q4_K_M --steps 128 --warmup 16 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('Q4_K_M split-V tg', d['value'], 'tok/s', d['ms_per_step'], 'ms  prefill', d['prefill_tok_s'])"
python bench.py --ftype Q4_K_M --steps 128 --warmup 16 2>/dev/null | grep '^{' > gpurun_out/bench_Q4_K_M.json; python -c "import json; d=json.load(open('gpurun_out/bench_Q4_K_M.json')); print('Q4_K_M tg', d['value'], 'tok/s', d['ms_per_step'], 'ms  prefill', d['prefill_tok_s'], 'e2e', d['e2e']['value'])"
python bench.py --steps 128 --warmup 16 2>/dev/null | grep '^{' > gpurun_out/bench_F16.json; python -c "import json; d=json.load(open('gpurun_out/bench_F16.json')); print('F16 tg', d['value'], 'tok/s', d['ms_per_step'], 'ms  prefill', d['prefill_tok_s'], 'e2e', d['e2e']['value'], d['clocks'])"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
