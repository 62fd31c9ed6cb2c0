#!/usr/bin/env bash
# full-set capture of the hot kernels (decode mat-vec, tcgen05 GEMM, decode attention); ONE ncu invocation
mkdir -p gpurun_out
FT=${1:-F16}
CMD="python tools/prof_step.py --ftype $FT --layers 2 --prefill 512 --decode 2"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/prof_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'gemv_dense_v3|gemm_tcgen05|attn_decode_rope|attn_prefill' -c 20 -o gpurun_out/prof_hot_$FT -f $CMD > gpurun_out/ncu_hot.log 2>&1
echo "full rc=$?"
ls -la gpurun_out/*.ncu-rep
