#!/usr/bin/env bash
# BASELINE.json configs 4 and 5, both arms, on one multi-GPU box.  Results land in gpurun_out/ as JSON lines.
#   config 4: Llama-3-70B Q4_K_M, 2 pipeline stages x TP=4, batched decode B = 1 / 8 / 32
#   config 5: Mixtral-8x7B Q4_K_M, PP x TP, concurrent requests through the orchestrator (aggregate tok/s, TTFT p50)
# usage: tools/run_baseline_configs.sh <pp> <tp> [model4 preset] [model5 preset] [n_layer override for a dry run] [arms: ours,ref]
set -uo pipefail
PP=${1:-2}; TP=${2:-4}; M4=${3:-llama-3-70b}; M5=${4:-mixtral-8x7b}; NL=${5:-0}; ARMS=${6:-ours,ref}
N=$((PP * TP))
MD=${DLP_MODEL_DIR:-/tmp/dlp_models}; export DLP_MODEL_DIR=$MD
OUT=gpurun_out; mkdir -p $OUT $MD
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
REF=baseline/_ref/bin
python baseline/reference.py > $OUT/cfg_ref_setup.log 2>&1
export LD_LIBRARY_PATH=$REF:${LD_LIBRARY_PATH:-}
GPUS=$(seq -s, 0 $((N - 1)))

gen() {  # preset ftype -> path
  python - "$1" "$2" "$NL" <<'EOF'
import os, sys, time
sys.path.insert(0, os.getcwd())
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
from distributed_llm_pipeline_b200.models.config import preset
name, ftype, nl = sys.argv[1], sys.argv[2], int(sys.argv[3])
path = f"{os.environ['DLP_MODEL_DIR']}/{name}{'-%dL' % nl if nl else ''}-{ftype}.gguf"
if not os.path.exists(path):
    t = time.time()
    write_synthetic_gguf(path + ".tmp", preset(name, **({"n_layer": nl} if nl else {})), ftype, seed=0)
    os.replace(path + ".tmp", path)
    print(f"wrote {path}: {os.path.getsize(path) / 1e9:.1f} GB in {time.time() - t:.0f} s", file=sys.stderr)
print(path)
EOF
}

clocks() { nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv,noheader -lms 500 > "$1" & echo $!; }

ONLY=${ONLY:-45}
# ---------------- config 4 ----------------
if [[ $ONLY == *4* ]]; then
P4=$(gen "$M4" Q4_K_M)
if [[ $ARMS == *ours* ]]; then
  C=$(clocks $OUT/cfg4_ours_clocks.csv)
  timeout 300 $TR tools/dlp_bench.py -m "$P4" --pp $PP --tp $TP -p 512 -n 32 -r 2 --batch 1,8,32 -c 1024 -o jsonl > $OUT/cfg4_ours.jsonl 2> $OUT/cfg4_ours.err
  RC=$?; echo "cfg4 ours rc=$RC"
  if [[ $RC != 0 ]]; then   # retry without graph capture of the sequence-parallel batched path
    tail -3 $OUT/cfg4_ours.err | cut -c1-300
    DLP_SP_GRAPH=0 timeout 300 $TR tools/dlp_bench.py -m "$P4" --pp $PP --tp $TP -p 512 -n 32 -r 2 --batch 1,8,32 -c 1024 -o jsonl > $OUT/cfg4_ours.jsonl 2> $OUT/cfg4_ours_retry.err
    echo "cfg4 ours (DLP_SP_GRAPH=0) rc=$?"
  fi
  kill $C 2>/dev/null
  cat $OUT/cfg4_ours.jsonl | cut -c1-400
fi
if [[ $ARMS == *ref* ]]; then
  MODES="layer row"; [[ $ARMS == *ref-layer* ]] && MODES="layer"
  for sm in $MODES; do
    C=$(clocks $OUT/cfg4_ref_${sm}_clocks.csv)
    CUDA_VISIBLE_DEVICES=$GPUS timeout 200 $REF/llama-batched-bench -m "$P4" -c 8192 -b 2048 -ub 512 -npp 16 -ntg 32 -npl 1,8,32 -ngl 99 -fa 1 -sm $sm --output-format jsonl \
      > $OUT/cfg4_ref_$sm.jsonl 2> $OUT/cfg4_ref_$sm.err
    echo "cfg4 ref -sm $sm rc=$?"; kill $C 2>/dev/null
    grep -h '{' $OUT/cfg4_ref_$sm.jsonl $OUT/cfg4_ref_$sm.err | grep n_kv | cut -c1-300
  done
fi
fi

# ---------------- config 5 ----------------
if [[ $ONLY == *5* ]]; then
P5=$(gen "$M5" Q4_K_M)
VOC=32000
wait_port() {  # until GET /health answers 200 (llama-server binds its port before the model is loaded)
  python - "$1" <<'EOF2'
import socket, sys, time
port = int(sys.argv[1])
for _ in range(300):
    try:
        s = socket.create_connection(("127.0.0.1", port), timeout=2)
        s.sendall(b"GET /health HTTP/1.1\r\nHost: x\r\nConnection: close\r\n\r\n")
        d = s.recv(4096)
        s.close()
        if b" 200 " in d.split(b"\r\n", 1)[0]:
            sys.exit(0)
    except OSError:
        pass
    time.sleep(0.5)
sys.exit(1)
EOF2
}
if [[ $ARMS == *ours* ]]; then
  C=$(clocks $OUT/cfg5_ours_clocks.csv)
  DLP_MASTER_PORT=29621 setsid python -m distributed_llm_pipeline_b200.serving.orchestrator -m "$P5" --pp $PP --tp $TP -np 16 -c 1024 -b 2048 --port 3105 -n 64 \
      > $OUT/cfg5_ours_server.log 2>&1 &
  SRV=$!
  if wait_port 3105; then
    python tools/serve_bench.py --port 3105 --route /completion -c 1,4,16 --prompt-tokens 128 -n 64 --vocab $VOC --tag ours-completion -o $OUT/cfg5_ours_completion.json
    python tools/serve_bench.py --port 3105 --route /chat -c 1,4,16 --prompt-tokens 128 -n 64 --vocab $VOC --tag ours-chat -o $OUT/cfg5_ours_chat.json
    grep -h "warm-up\|grid:" $OUT/cfg5_ours_server.log | head -3
  else echo "cfg5 ours: server did not come up"; grep -v "Warn\|warn\|from_numpy" $OUT/cfg5_ours_server.log | tail -8 | cut -c1-300; fi
  kill -INT -- -$SRV 2>/dev/null; sleep 3; kill -KILL -- -$SRV 2>/dev/null; kill $C 2>/dev/null
fi
if [[ $ARMS == *ref* ]]; then
  C=$(clocks $OUT/cfg5_ref_clocks.csv)
  CUDA_VISIBLE_DEVICES=$GPUS setsid $REF/llama-server -m "$P5" -np 16 -cb -sm layer -ngl 99 -fa 1 -c 16384 -b 2048 --host 127.0.0.1 --port 3106 > $OUT/cfg5_ref_server.log 2>&1 &
  SRV=$!
  if wait_port 3106; then
    sleep 2
    python tools/serve_bench.py --port 3106 --route /completion -c 1,4,16 --prompt-tokens 128 -n 64 --vocab $VOC --tag ref-llama-server -o $OUT/cfg5_ref_completion.json
  else echo "cfg5 ref: llama-server did not come up"; tail -5 $OUT/cfg5_ref_server.log; fi
  kill -INT -- -$SRV 2>/dev/null; sleep 2; kill -KILL -- -$SRV 2>/dev/null; kill $C 2>/dev/null
fi
fi
echo "configs done"
