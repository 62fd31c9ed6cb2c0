#!/usr/bin/env bash
# full-set capture of the block-quant decode mat-vec; ONE ncu invocation
mkdir -p gpurun_out
FT=${1:-Q4_K_M}
CMD="python tools/prof_step.py --ftype $FT --layers 2 --prefill 16 --decode 2"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/prof_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'gemv_quant_v3' -s 8 -c 5 -o gpurun_out/prof_quant_$FT -f $CMD > gpurun_out/ncu_quant.log 2>&1
echo "full rc=$?"
ls -la gpurun_out/prof_quant_$FT.ncu-rep
