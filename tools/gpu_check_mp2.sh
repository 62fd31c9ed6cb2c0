#!/usr/bin/env bash
# multi-GPU sanity after engine changes: parallel tests + fused-TP bench (default steps)
N=${1:-2}
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_parallel_gpu.py -x -q --timeout 250 -p no:cacheprovider 2>&1 | tail -5
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29801 bench.py --gpus $N --steps 128 --warmup 16 --tp-impl fused 2>/dev/null | grep '^{' > gpurun_out/bench_tp$N.json
python -c "import json; d=json.load(open('gpurun_out/bench_tp$N.json')); print('TP$N fused tg', d['value'], 'tok/s', d['ms_per_step'], 'ms prefill', d['prefill_tok_s'], 'e2e', d['e2e']['value'])"
