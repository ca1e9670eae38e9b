#!/bin/bash
# fused spatial kernel: timing probe, then one ncu capture of it (source counters + warp states)
mkdir -p gpurun_out
timeout 200 python tools/spatial_probe.py 128 10 > gpurun_out/spatial_probe.txt 2>&1 || { tail -5 gpurun_out/spatial_probe.txt; exit 1; }
cat gpurun_out/spatial_probe.txt
timeout 400 ncu --set full --import-source on --clock-control none -k regex:spatial_fused --launch-skip 6 -c 2 -o gpurun_out/spatial_fused_r2 -f python tools/spatial_probe.py 128 4 > gpurun_out/spatial_ncu.log 2>&1
tail -3 gpurun_out/spatial_ncu.log
