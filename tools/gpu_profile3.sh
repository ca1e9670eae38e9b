#!/bin/bash
set -u
mkdir -p gpurun_out
CMD="python bench.py --only-train --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain4.log 2>&1 &&
timeout 1700 ncu --set full --clock-control none --import-source on -k regex:"wgrad_umma|agg_umma_kernel|agg_dA_umma" -s 165 -c 55 -o gpurun_out/prof_wgrad_agg $CMD > gpurun_out/ncu_wgrad_agg.log 2>&1
echo "rc=$?" >> gpurun_out/ncu_wgrad_agg.log
tail -n 4 gpurun_out/ncu_wgrad_agg.log
