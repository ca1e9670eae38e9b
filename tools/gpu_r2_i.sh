#!/bin/bash
# A/B of the fused spatial kernel's epilogue split on the inference workload (same box, alternating), plus its parity tests
for i in 1 2 3; do for sub in 1 2; do
  STGCN_SPATIAL_SUB=$sub timeout 120 python bench.py --workload cfg3 --steps 20 --warmup 5 --no-cpu-baseline --no-torch-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('SUB=$sub', round(d['value']), round(d['ms_per_step'],3), d['kernel_ms_per_step'].get('umma_spatial'))"
done; done
STGCN_SPATIAL_SUB=2 timeout 200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_block.py tests/test_gpu_model.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -1
