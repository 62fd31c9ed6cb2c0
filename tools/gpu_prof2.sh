#!/usr/bin/env bash
# launch list only (device time + dram bytes per kernel) for a decode-heavy run
FT=${1:-Q8_0}
mkdir -p gpurun_out
CMD="python tools/prof_step.py --ftype $FT --layers 3 --prefill 0 --decode 3"
$CMD > gpurun_out/prof_plain.log 2>&1 || { tail -20 gpurun_out/prof_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,sm__inst_executed.sum,launch__grid_size,launch__registers_per_thread --clock-control none --csv --log-file gpurun_out/launches_$FT.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "rc=$?"
