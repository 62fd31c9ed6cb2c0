#!/usr/bin/env bash
# ncu captures (1 GPU): launch list with device times, then full sets of the top kernels. Results -> gpurun_out/
mkdir -p gpurun_out
FT=${1:-F16}
CMD="python tools/prof_step.py --ftype $FT --layers 4 --prefill 512 --decode 3"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/prof_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_$FT.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemv_ -c 8 -o gpurun_out/prof_gemv_$FT -f $CMD > gpurun_out/ncu_gemv.log 2>&1
echo "gemv full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -c 6 -o gpurun_out/prof_gemm_$FT -f $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm full rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_$FT.csv
