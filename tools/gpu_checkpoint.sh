#!/bin/bash
# Round checkpoint: full bench line, ncu launch list, DRAM traffic of every conv_gemm launch of one step, full capture of a few launches.
# The ncu passes profile the EAGER step (kernel by kernel); the bench line itself replays the captured CUDA graph.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
CMD="python bench.py --only-train --eager --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 python bench.py > gpurun_out/bench_full.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench_full.log
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?" >> gpurun_out/ncu_list.log
# one training step launches 39-40 conv_gemm kernels (10 mixing + 10 temporal forward, 10 temporal + 9-10 mixing data gradients);
# 6 warm-up steps precede the timed one; any 40 consecutive launches cover one whole step
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:conv_gemm_umma -s 240 -c 40 --csv --log-file gpurun_out/conv_gemm_traffic.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
echo "traffic rc=$?" >> gpurun_out/ncu_traffic.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_umma -s 240 -c 8 -o gpurun_out/prof_conv_gemm $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?" >> gpurun_out/ncu_full.log
tail -n 2 gpurun_out/bench_full.log gpurun_out/ncu_list.log gpurun_out/ncu_traffic.log gpurun_out/ncu_full.log | cut -c1-300
