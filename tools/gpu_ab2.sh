#!/usr/bin/env bash
# A/B of decode knobs: each line "ENV=val ..." is one bench run (F16, 64 steps)
mkdir -p gpurun_out
python -m pytest tests/test_kernels_gpu.py -q -k "gemv or fused_allreduce or moe" -p no:cacheprovider 2>&1 | tail -1
while read -r envs; do
  [ -z "$envs" ] && continue
  env $envs timeout 300 python bench.py --steps 64 --warmup 8 ${BENCH_EXTRA:-} 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$envs', '->', d['value'], 'tok/s', d['ms_per_step'], 'ms  prefill', d['prefill_tok_s'])"
done <<'CFG'
DLP_GEMV_UD=42
DLP_GEMV_UD=24
CFG
