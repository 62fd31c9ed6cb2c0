#!/usr/bin/env bash
mkdir -p gpurun_out
for pdl in 1 0; do
  DLP_PDL=$pdl timeout 400 python bench.py --steps 48 --warmup 8 > gpurun_out/bench_pdl$pdl.log 2>&1
  grep '^{' gpurun_out/bench_pdl$pdl.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('PDL=$pdl', d['value'], 'tok/s', d['ms_per_step'], 'ms', 'prefill', d['prefill_tok_s'], 'ttft', d['ttft_ms_p50'])"
done
timeout 900 bash tools/gpu_prof.sh F16 2>&1 | tail -8
