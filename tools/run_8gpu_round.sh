#!/usr/bin/env bash
# One 8-GPU session, most important results first (box time is charged x8): tp8 correctness test, headline bench at N=8,
# BASELINE configs 4 and 5 (ours, then the reference), device timeline at tp8.  Everything lands in gpurun_out/.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT /tmp/dlp_models
T0=$(date +%s); el() { echo "[t+$(( $(date +%s) - T0 ))s] $*"; }
# the two big synthetic checkpoints are written in the background while the first steps run
( python - <<'PY' > $OUT/r8_gen.log 2>&1
import os, sys, time
sys.path.insert(0, os.getcwd())
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
from distributed_llm_pipeline_b200.models.config import preset
for name in ("llama-3-70b", "mixtral-8x7b"):
    path = f"/tmp/dlp_models/{name}-Q4_K_M.gguf"
    if not os.path.exists(path):
        t = time.time()
        write_synthetic_gguf(path + ".tmp", preset(name), "Q4_K_M", seed=0)
        os.replace(path + ".tmp", path)
        print(f"wrote {path}: {os.path.getsize(path) / 1e9:.1f} GB in {time.time() - t:.0f} s", flush=True)
PY
) &
GEN=$!
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811"
timeout 150 python -m pytest tests/test_parallel_gpu.py -x -q -k "tp_wide" --timeout 120 > $OUT/r8_pytest.log 2>&1; el "tp_wide tests rc=$? $(tail -1 $OUT/r8_pytest.log)"
timeout 200 $TR bench.py --gpus 8 --steps 20 --warmup 5 > $OUT/r8_bench8_fused.json 2> $OUT/r8_bench8_fused.err; el "bench8 rc=$?"
grep '^{' $OUT/r8_bench8_fused.json | cut -c1-1700
timeout 120 $TR tools/timeline_tp.py --out $OUT/r8_tl_tp8.json > $OUT/r8_tl_tp8.txt 2>&1; el "timeline rc=$?"; grep -v "Warn\|warn\|from_numpy" $OUT/r8_tl_tp8.txt | tail -14
wait $GEN; el "checkpoints: $(cat $OUT/r8_gen.log | tr '\n' ' ')"
timeout 420 tools/run_baseline_configs.sh 2 4 llama-3-70b mixtral-8x7b 0 ours 2>&1 | grep -v "^NCCL\|Warn\|warn" | cut -c1-420; el "configs ours done"
timeout 420 tools/run_baseline_configs.sh 2 4 llama-3-70b mixtral-8x7b 0 ref 2>&1 | cut -c1-330; el "configs ref done"
