#!/bin/bash
# ncu --set full of the transposed aggregation (one whole step), 4 forward conv launches (C=64 layers) and the last 4 wgrad launches.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
CMD="python bench.py --only-train --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:agg_umma_kernel -s 30 -c 10 -o gpurun_out/prof_aggbwd $CMD > gpurun_out/ncu4a.log 2>&1
echo "rc=$?" >> gpurun_out/ncu4a.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_umma -s 90 -c 4 -o gpurun_out/prof_conv64 $CMD > gpurun_out/ncu4b.log 2>&1
echo "rc=$?" >> gpurun_out/ncu4b.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wgrad_umma_kernel -s 76 -c 4 -o gpurun_out/prof_wgrad64 $CMD > gpurun_out/ncu4c.log 2>&1
echo "rc=$?" >> gpurun_out/ncu4c.log
tail -n 2 gpurun_out/ncu4a.log gpurun_out/ncu4b.log gpurun_out/ncu4c.log; ls -la gpurun_out/*.ncu-rep
