#!/bin/bash
# A/B of one environment switch (default: STGCN_NO_HALO) after the parity tests:  gpu_ab.sh [ENV_VAR]
set -u
VAR=${1:-STGCN_NO_HALO}
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/ab_tests.log
for mode in default switched; do
  if [ $mode = switched ]; then export $VAR=1; else unset $VAR; fi
  timeout 300 python bench.py --only-train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab_train_$mode.log 2>&1
  timeout 300 python bench.py --only-infer --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab_infer_$mode.log 2>&1
done
tail -n 12 gpurun_out/ab_tests.log
for f in gpurun_out/ab_train_*.log gpurun_out/ab_infer_*.log; do echo $f; tail -n 1 $f | cut -c1-1500; done
