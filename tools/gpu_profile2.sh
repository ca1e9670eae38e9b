#!/bin/bash
# ncu section captures (one step) of the non-GEMM-engine kernels: wgrad, aggregation, dA, BN passes
set -u
mkdir -p gpurun_out
CMD="python bench.py --only-train --steps 1 --warmup 3 --no-cpu-baseline"
SECS="--section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section SchedulerStats --section LaunchStats --section Occupancy --section ComputeWorkloadAnalysis"
timeout 300 $CMD > gpurun_out/plain3.log 2>&1 &&
timeout 1500 ncu $SECS --clock-control none --import-source on -k regex:"wgrad_umma|agg_dA_umma|agg_umma_kernel|colreduce|bn_bwd_apply|affine_act" -s 519 -c 173 -o gpurun_out/prof_others $CMD > gpurun_out/ncu_others.log 2>&1
echo "others rc=$?" >> gpurun_out/ncu_others.log
tail -n 4 gpurun_out/ncu_others.log
