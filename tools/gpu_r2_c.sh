#!/bin/bash
mkdir -p gpurun_out
timeout 400 python tools/bench_block.py --out gpurun_out/block_sweep.json > gpurun_out/block_sweep.log 2>&1; echo "sweep rc=$?"; tail -2 gpurun_out/block_sweep.log | cut -c1-300
CMD="python bench.py --only-train --eager --steps 2 --warmup 3"
timeout 200 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu.log | cut -c1-300; wc -l gpurun_out/launches.csv
