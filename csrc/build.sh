#!/usr/bin/env bash
# Builds libdlp_b200.so (sm_100a only) in-tree: distributed_llm_pipeline_b200/ops/libdlp_b200.so
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../distributed_llm_pipeline_b200/ops"
OBJ="$HERE/build"
mkdir -p "$OBJ" "$OUT"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --use_fast_math -Xptxas -v)
pids=()
for f in "$HERE"/*.cu; do
  b="$(basename "$f" .cu)"
  if [[ ! -f "$OBJ/$b.o" || "$f" -nt "$OBJ/$b.o" || -n "$(find "$HERE" -name '*.cuh' -newer "$OBJ/$b.o" -print -quit)" || "$HERE/../include/dlp_b200.h" -nt "$OBJ/$b.o" ]]; then
    ( "$NVCC" "${FLAGS[@]}" -c "$f" -o "$OBJ/$b.o" > "$OBJ/$b.log" 2>&1 || { cat "$OBJ/$b.log"; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [[ -n "$p" ]] && wait "$p"; done
"$NVCC" -shared -o "$OUT/libdlp_b200.so" "$OBJ"/*.o -lcudart_static -lpthread -ldl -lrt
echo "built $OUT/libdlp_b200.so"
