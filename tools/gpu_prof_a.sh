#!/usr/bin/env bash
# launch list (device time + DRAM bytes per kernel) of pp512 + 3 eager decode steps, 4 layers; ONE ncu invocation
mkdir -p gpurun_out
FT=${1:-F16}
CMD="python tools/prof_step.py --ftype $FT --layers 4 --prefill 512 --decode 3"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/prof_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread --clock-control none --csv --log-file gpurun_out/launches_$FT.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
grep -c '^"' gpurun_out/launches_$FT.csv
