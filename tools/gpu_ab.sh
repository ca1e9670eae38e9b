#!/bin/bash
# A/B of the fused spatial kernel against the aggregation + GEMM pair, after the parity tests.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/ab_tests.log
for mode in fused unfused; do
  if [ $mode = fused ]; then export STGCN_FUSED_SPATIAL=1; else unset STGCN_FUSED_SPATIAL; fi
  timeout 300 python bench.py --only-train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab_train_$mode.log 2>&1
  timeout 300 python bench.py --only-infer --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab_infer_$mode.log 2>&1
done
tail -n 3 gpurun_out/ab_tests.log
for f in gpurun_out/ab_train_*.log gpurun_out/ab_infer_*.log; do echo $f; tail -n 1 $f | cut -c1-1500; done
