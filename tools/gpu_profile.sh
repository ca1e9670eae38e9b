#!/bin/bash
# One gpurun call: smoke, plain bench leg, ncu launch list of the same command, ncu --set full of the top kernel.
set -u
mkdir -p gpurun_out
CMD="python bench.py --only-train --steps 2 --warmup 3 --no-cpu-baseline"
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?" >> gpurun_out/ncu_list.log
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_umma -s 120 -c 8 -o gpurun_out/prof_conv_gemm $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?" >> gpurun_out/ncu_full.log
tail -n 3 gpurun_out/smoke.log gpurun_out/plain.log gpurun_out/ncu_list.log gpurun_out/ncu_full.log
