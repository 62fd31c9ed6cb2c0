#!/usr/bin/env bash
mkdir -p gpurun_out
python bench.py --steps 64 --warmup 8 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('F16 tg', d['value'], 'tok/s', d['ms_per_step'], 'ms  prefill', d['prefill_tok_s'])"
M=/tmp/dlp_models/llama-3-8b-F16.gguf
for B in 8 32; do
  timeout 200 python tools/dlp_bench.py -m $M -p 0 -n 64 -r 2 --batch $B -c 1024 -o csv 2>/dev/null | tail -1
done
LD_LIBRARY_PATH=baseline/_ref/bin timeout 300 baseline/_ref/bin/llama-batched-bench -m $M -c 8192 -b 2048 -ub 512 -npp 16 -ntg 64 -npl 8,32 -ngl 99 -fa 1 2>/dev/null | grep -E "^\|" | tail -4
