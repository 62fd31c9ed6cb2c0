#!/usr/bin/env bash
# new-architecture kernels/engine tests + A/B of the L2 weight prefetch in the RoPE kernel
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -k "rope" tests/test_engine_gpu.py -x -q 2>&1 | tail -6
for PF in 1 0; do
  DLP_L2_PREFETCH=$PF python bench.py --steps 64 --warmup 8 2>/dev/null | grep '^{' > gpurun_out/bench_pf$PF.json
  python -c "import sys,json; d=json.load(open('gpurun_out/bench_pf$PF.json')); print('L2 prefetch=$PF  F16 tg', d['value'], 'tok/s', d['ms_per_step'], 'ms  prefill', d['prefill_tok_s'], 'e2e', d['e2e']['value'])"
done
for Q in Q8_0 Q4_K_M; do for PF in 1 0; do
  DLP_L2_PREFETCH=$PF python bench.py --ftype $Q --steps 64 --warmup 8 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$Q prefetch=$PF tg', d['value'], 'tok/s  prefill', d['prefill_tok_s'])"
done; done
