#!/usr/bin/env bash
# Decode and prefill rate with 0 / 4096 / 32768 tokens already in the KV cache (llama-bench -d), both arms, one GPU.
# usage: tools/run_depth_sweep.sh [model preset] [ftype] [depths]
set -uo pipefail
M=${1:-llama-3-8b}; FT=${2:-F16}; D=${3:-0,4096,32768}
MD=${DLP_MODEL_DIR:-/tmp/dlp_models}; export DLP_MODEL_DIR=$MD
OUT=gpurun_out; mkdir -p $OUT $MD
P=$(python - "$M" "$FT" <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
from distributed_llm_pipeline_b200.models.config import preset
name, ftype = sys.argv[1], sys.argv[2]
path = f"{os.environ['DLP_MODEL_DIR']}/{name}-{ftype}.gguf"
if not os.path.exists(path):
    write_synthetic_gguf(path + ".tmp", preset(name), ftype, seed=0)
    os.replace(path + ".tmp", path)
print(path)
PY
)
echo "model $P"
timeout 120 python tools/dlp_bench.py -m "$P" -p 512 -n 32 -d "$D" -r 1 -ub 2048 -o jsonl > $OUT/depth_ours.jsonl 2> $OUT/depth_ours.err
echo "ours rc=$?"; cut -c1-260 $OUT/depth_ours.jsonl; tail -2 $OUT/depth_ours.err | cut -c1-300
python baseline/reference.py > $OUT/depth_ref_setup.log 2>&1
LD_LIBRARY_PATH=baseline/_ref/bin:${LD_LIBRARY_PATH:-} timeout 120 baseline/_ref/bin/llama-bench -m "$P" -p 512 -n 32 -d "$D" -r 1 -ngl 99 -fa 1 -o jsonl > $OUT/depth_ref.jsonl 2> $OUT/depth_ref.err
echo "ref rc=$?"; python - <<'PY'
import json
for l in open("gpurun_out/depth_ref.jsonl"):
    try:
        d = json.loads(l)
    except ValueError:
        continue
    print("ref", "pp" if d.get("n_prompt") else "tg", d.get("n_prompt") or d.get("n_gen"), "depth", d.get("n_depth"), round(d.get("avg_ts", 0), 1))
PY
