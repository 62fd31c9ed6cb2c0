#!/usr/bin/env bash
mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0,'.')
from bench import ensure_model
print(ensure_model('llama-3-8b','Q8_0',0,lambda:None))
PY
M=/tmp/dlp_models/llama-3-8b-Q8_0.gguf
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29821 tools/dlp_bench.py -m $M --pp 2 --tp 1 -p 512 -n 128 -r 3 -o json > gpurun_out/pp2_ours.json 2> gpurun_out/pp2_ours.err; echo "ours rc=$?"; grep -E '"test"|"avg_ts"' gpurun_out/pp2_ours.json; grep -E "Error|error" gpurun_out/pp2_ours.err | head -5
