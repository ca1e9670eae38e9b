#!/bin/bash
# round-2 final launch lists (ncu gpu__time_duration, --clock-control none) of the training step and of the cfg3 inference forward,
# each after the same command exited 0 without ncu
set -u
mkdir -p gpurun_out
CMD="python bench.py --only-train --eager --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_train.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_train_final.csv $CMD > gpurun_out/ncu_list_train.log 2>&1
echo "train list rc=$?"
CMD2="python bench.py --workload cfg3 --only-infer --eager --steps 2 --warmup 3 --no-cpu-baseline --no-torch-baseline"
timeout 300 $CMD2 > gpurun_out/plain_infer.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_infer_final.csv $CMD2 > gpurun_out/ncu_list_infer.log 2>&1
echo "infer list rc=$?"
tail -n 1 gpurun_out/plain_train.log | cut -c1-200; tail -n 1 gpurun_out/plain_infer.log | cut -c1-200
