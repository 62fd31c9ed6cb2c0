#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -k "prefill" -x -q 2>&1 | tail -3
timeout 600 python -m pytest tests/test_engine_gpu.py -x -q 2>&1 | tail -3
python bench.py --steps 32 --warmup 4 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('F16 tg', d['value'], 'tok/s  prefill', d['prefill_tok_s'])"
python bench.py --ftype Q8_0 --steps 32 --warmup 4 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('Q8_0 tg', d['value'], 'tok/s  prefill', d['prefill_tok_s'])"
