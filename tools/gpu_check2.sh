#!/usr/bin/env bash
mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 to=$2; shift 2; echo "=== $name ===" | tee -a gpurun_out/summary.txt
  timeout "$to" "$@" > "gpurun_out/$name.log" 2>&1; local rc=$?
  echo "$name rc=$rc" | tee -a gpurun_out/summary.txt; tail -n 12 "gpurun_out/$name.log"; }
run zz_alone 300 python -m pytest tests/test_zz_no_alloc.py -q --timeout 120 -p no:cacheprovider
run all_gpu 900 python -m pytest tests -m gpu -x -q --timeout 180 -p no:cacheprovider
run smoke 300 python __graft_entry__.py smoke
run bench 1200 python bench.py --steps 64 --warmup 8
run bench_ref 1500 python bench.py --impl reference --steps 64 --warmup 8
cat gpurun_out/summary.txt
