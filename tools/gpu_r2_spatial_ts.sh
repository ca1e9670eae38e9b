#!/bin/bash
# tensor-memory variant of the fused spatial kernel: parity tests that run it, then the timing probe and the segment timers
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_block.py tests/test_gpu_fullsize.py tests/test_gpu_train.py tests/test_gpu_model.py -q -x -k "inference or prepared or captured_forward or eval or full_size" 2>&1 | tail -5
echo "--- probe"; timeout 120 python tools/spatial_probe.py 128 10 2>&1 | tail -6
echo "--- probe ${PROBE2_ENV}"; env ${PROBE2_ENV} timeout 120 python tools/spatial_probe.py 128 10 2>&1 | tail -6
timeout 150 python tools/spatial_trace.py 2>&1 | grep "sp trace" | awk "NR%2==0"
