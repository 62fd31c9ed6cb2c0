#!/usr/bin/env bash
# compute-sanitizer over a subset of the kernel tests (the reference has no compute-sanitizer integration).
# NOTE: on the development pool compute-sanitizer is disabled ("closed on this pool"): this script is for boxes where it is
# available; in-tree the kernels are guarded by DLP_REQUIRE argument checks, small-shape tests and the CPU-oracle comparison.
#   tools/sanitize_gpu.sh memcheck|racecheck|synccheck ["pytest -k expression"]
TOOL=${1:-memcheck}
EXPR=${2:-"onepass and 128-8-2 or gemv_store_and_add and 1- or rope_kv_write"}
mkdir -p gpurun_out
timeout ${SAN_TIMEOUT:-200} compute-sanitizer --tool $TOOL --error-exitcode 7 --print-limit 20 \
  python -m pytest tests/test_kernels_gpu.py -x -q -k "$EXPR" -p no:cacheprovider > gpurun_out/sanitize_$TOOL.log 2>&1
rc=$?
echo "compute-sanitizer $TOOL rc=$rc"
grep -E "passed|failed|ERROR SUMMARY|========= (Invalid|Race|Barrier|Error)" gpurun_out/sanitize_$TOOL.log | head -20
exit 0
