#!/usr/bin/env bash
# Round-trip GPU check used through gpurun: kernel numerics, engine tests, smoke, short bench.
# Each stage runs in its own process under a timeout so a faulting kernel cannot take the others down.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, timeout, cmd...
  local name=$1 to=$2; shift 2
  echo "=== $name ===" | tee -a gpurun_out/summary.txt
  timeout "$to" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "$name rc=$rc" | tee -a gpurun_out/summary.txt
  tail -n 25 "gpurun_out/$name.log"
}
: > gpurun_out/summary.txt
run kernels 900 python -m pytest tests/test_kernels_gpu.py -q --timeout 180 -p no:cacheprovider
run engine 600 python -m pytest tests/test_engine_gpu.py -q --timeout 180 -p no:cacheprovider
run smoke 300 python __graft_entry__.py smoke
if [[ "${1:-}" != "nobench" ]]; then
  run bench 1200 python bench.py --steps "${BENCH_STEPS:-64}" --warmup 8 ${BENCH_ARGS:-}
fi
cat gpurun_out/summary.txt
