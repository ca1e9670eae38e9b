#!/bin/bash
# ncu --set full of the implicit-GEMM engine at given CxN shapes of tools/bench_conv.py:  gpu_profile_conv.sh 64x192 ...
set -u
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
python tools/bench_conv.py "$@" > gpurun_out/conv_plain.log 2>&1 &&
ITERS=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_umma -o gpurun_out/prof_convshape python tools/bench_conv.py "$@" > gpurun_out/ncu_conv.log 2>&1
echo "rc=$?" >> gpurun_out/ncu_conv.log
cat gpurun_out/conv_plain.log; tail -n 2 gpurun_out/ncu_conv.log
