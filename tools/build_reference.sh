#!/usr/bin/env bash
# Builds the UNMODIFIED reference (/root/reference/llama.cpp) with its own CMake for sm_100 and installs the
# binaries the reference arm needs into baseline/_ref/bin.  Offline; nothing of this repository is on that path.
# usage: tools/build_reference.sh [build_dir]   (default /tmp/refbuild)
#        tools/build_reference.sh cpu-tools [build_dir]   CPU-only static builds of llama-quantize / llama-imatrix / llama-gguf-split /
#                                                         llama-export-lora / llama-cvector-generator / llama-tokenize into baseline/_ref/bin_cpu
#                                                         (what tests/test_tools_vs_reference.py exchanges files with; about 4 minutes on 8 cores)
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${DLP_REFERENCE_SRC:-/root/reference/llama.cpp}"
BLD="${1:-/tmp/refbuild}"
OUT="$REPO/baseline/_ref/bin"
JOBS="${DLP_REFERENCE_JOBS:-$(nproc)}"
[ -d "$SRC" ] || { echo "reference source $SRC not found" >&2; exit 3; }
if [ "${1:-}" = "cpu-tools" ]; then
  BLD="${2:-/tmp/refcpu}"; OUT="$REPO/baseline/_ref/bin_cpu"; mkdir -p "$BLD" "$OUT"
  TOOLS="llama-quantize llama-imatrix llama-gguf-split llama-export-lora llama-cvector-generator llama-tokenize"
  cmake -S "$SRC" -B "$BLD" -DCMAKE_BUILD_TYPE=Release -DGGML_CUDA=OFF -DGGML_NATIVE=OFF -DGGML_AVX2=ON -DGGML_FMA=ON -DGGML_F16C=ON \
    -DBUILD_SHARED_LIBS=OFF -DLLAMA_CURL=OFF -DLLAMA_BUILD_TESTS=OFF -DLLAMA_BUILD_EXAMPLES=OFF -DLLAMA_BUILD_TOOLS=ON -DLLAMA_BUILD_SERVER=OFF \
    > "$BLD/configure.log" 2>&1
  cmake --build "$BLD" -j "$JOBS" --target $TOOLS > "$BLD/build.log" 2>&1
  for f in $TOOLS; do cp -f "$BLD/bin/$f" "$OUT/"; done
  echo "reference CPU tools installed in $OUT"
  exit 0
fi
mkdir -p "$BLD" "$OUT"
cmake -S "$SRC" -B "$BLD" -G Ninja -DCMAKE_BUILD_TYPE=Release -DGGML_CUDA=ON -DGGML_RPC=ON \
  -DCMAKE_CUDA_ARCHITECTURES=100 -DGGML_NATIVE=OFF -DLLAMA_CURL=OFF -DLLAMA_BUILD_TESTS=OFF \
  -DLLAMA_BUILD_EXAMPLES=OFF -DLLAMA_BUILD_TOOLS=ON -DLLAMA_BUILD_SERVER=ON > "$BLD/configure.log" 2>&1
cmake --build "$BLD" -j "$JOBS" --target llama-bench llama-batched-bench llama-cli llama-server rpc-server llama-perplexity \
  > "$BLD/build.log" 2>&1
for f in llama-bench llama-batched-bench llama-cli llama-server rpc-server llama-perplexity; do
  cp -f "$BLD/bin/$f" "$OUT/"
done
find "$BLD/bin" -maxdepth 1 -name '*.so*' -exec cp -af {} "$OUT/" \;
strip --strip-unneeded "$OUT"/* 2>/dev/null || true
echo "reference installed in $OUT"
