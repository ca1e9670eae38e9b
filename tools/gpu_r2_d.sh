#!/bin/bash
# round 2 evidence: DRAM bytes + time of every launch of one training step, and one --set full capture of the top kernel
mkdir -p gpurun_out
CMD="python bench.py --only-train --eager --steps 1 --warmup 3"
timeout 200 $CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -3 gpurun_out/plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2100 --csv --log-file gpurun_out/step_dram.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "dram rc=$?"; wc -l gpurun_out/step_dram.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_umma -s 122 -c 6 -o gpurun_out/prof_conv_r2 -f $CMD > gpurun_out/ncu2.log 2>&1
echo "full rc=$?"; ls -la gpurun_out/prof_conv_r2.ncu-rep
