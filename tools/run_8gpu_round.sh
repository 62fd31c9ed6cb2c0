#!/usr/bin/env bash
# One 8-GPU session, most important results first (box time is charged x8): headline bench at N=8, device timeline at tp8,
# BASELINE config 4 (70B Q4_K_M, 2 x 4) on both arms, config 5 (Mixtral serving) on ours.  Checkpoints are written one at a
# time into the roomiest filesystem and deleted after use (a 42 GB + 28 GB + 16 GB set filled the scratch disk once and every
# later step died with ENOSPC).  Key results are echoed to stdout as well as saved under gpurun_out/.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
T0=$(date +%s); el() { echo "[t+$(( $(date +%s) - T0 ))s] $*"; }
best=/tmp; bestk=0
for d in /dev/shm /tmp "$PWD"; do k=$(df --output=avail -k "$d" 2>/dev/null | tail -1 | tr -d ' '); [[ -n "$k" && "$k" -gt "$bestk" ]] && { best=$d; bestk=$k; }; done
export DLP_MODEL_DIR=$best/dlp_models; mkdir -p $DLP_MODEL_DIR
el "model dir $DLP_MODEL_DIR ($((bestk / 1048576)) GB free); $(df -h /tmp /dev/shm 2>/dev/null | tail -2 | tr '\n' ';')"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811"
export TMPDIR=$best/tmp_$$; mkdir -p $TMPDIR

timeout 200 $TR bench.py --gpus 8 --steps 20 --warmup 5 > $OUT/r8_bench8_fused.json 2> $OUT/r8_bench8_fused.err; el "bench8 rc=$?"
grep '^{' $OUT/r8_bench8_fused.json | cut -c1-1700; tail -2 $OUT/r8_bench8_fused.err | cut -c1-300
timeout 120 $TR tools/timeline_tp.py --out $OUT/r8_tl_tp8.json > $OUT/r8_tl_tp8.txt 2>&1; el "timeline rc=$?"; grep -v "Warn\|warn\|from_numpy" $OUT/r8_tl_tp8.txt | tail -14
rm -f $DLP_MODEL_DIR/*.gguf

ONLY=4 timeout 400 tools/run_baseline_configs.sh 2 4 llama-3-70b mixtral-8x7b 0 "ours,ref-layer" 2>&1 | grep -v "^NCCL\|Warn\|warn" | cut -c1-420; el "config 4 done"
rm -f $DLP_MODEL_DIR/*.gguf
ONLY=5 timeout 300 tools/run_baseline_configs.sh 2 4 llama-3-70b mixtral-8x7b 0 ours 2>&1 | grep -v "^NCCL\|Warn\|warn" | cut -c1-420; el "config 5 done"
rm -rf $DLP_MODEL_DIR $TMPDIR
