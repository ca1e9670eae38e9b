#!/bin/bash
# plain run, then the ncu launch list (gpu__time_duration) of the same command
set -u
mkdir -p gpurun_out
CMD="python bench.py --only-train --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?" >> gpurun_out/ncu_list.log
tail -n 2 gpurun_out/plain.log gpurun_out/ncu_list.log
