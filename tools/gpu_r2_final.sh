#!/bin/bash
# round-2 closing run: full GPU suite, smoke, the default bench line, the reference arm, cfg3 / cfg5 lines, then one ncu capture of
# the tensor-memory spatial kernel (after the same command ran without ncu)
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py > gpurun_out/final_cfg4.json 2> gpurun_out/final_cfg4.err; tail -c 600 gpurun_out/final_cfg4.json; echo
timeout 300 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; tail -c 400 gpurun_out/final_ref.json; echo
timeout 300 python bench.py --workload cfg3 > gpurun_out/final_cfg3.json 2> gpurun_out/final_cfg3.err; python -c "import json; d=json.load(open('gpurun_out/final_cfg3.json')); print('cfg3', d['value'], d['ms_per_step'], d.get('e2e'), d.get('roofline_step', d.get('roofline')))"
timeout 300 python bench.py --workload cfg3 --dtype f32 --no-cpu-baseline > gpurun_out/final_cfg3_f32.json 2> gpurun_out/final_cfg3_f32.err; python -c "import json; d=json.load(open('gpurun_out/final_cfg3_f32.json')); print('cfg3 f32', d['value'], d['ms_per_step'])"
timeout 400 python bench.py --workload cfg5 --no-cpu-baseline > gpurun_out/final_cfg5.json 2> gpurun_out/final_cfg5.err; python -c "import json; d=json.load(open('gpurun_out/final_cfg5.json')); print('cfg5', d['value'], d['ms_per_step'], d.get('infer', {}).get('value'))"
timeout 100 python tools/spatial_probe.py 128 4 > /dev/null 2>&1 && timeout 300 ncu --set full --import-source on --clock-control none -k regex:spatial_fused_ts --launch-skip 4 -c 3 -o gpurun_out/spatial_ts_r2 -f python tools/spatial_probe.py 128 4 > gpurun_out/spatial_ts_ncu.log 2>&1
tail -2 gpurun_out/spatial_ts_ncu.log
