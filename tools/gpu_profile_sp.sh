#!/bin/bash
set -u
mkdir -p gpurun_out
CMD="python bench.py --only-infer --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/plain_sp.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spatial_fused -s 45 -c 9 -o gpurun_out/prof_spatial $CMD > gpurun_out/ncu_sp.log 2>&1
echo "rc=$?" >> gpurun_out/ncu_sp.log
tail -n 3 gpurun_out/ncu_sp.log
