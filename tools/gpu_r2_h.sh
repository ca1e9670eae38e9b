#!/bin/bash
N=${1:-8}; mkdir -p gpurun_out
python -c "
import sys; sys.path.insert(0,'.')
from stgcn_b200 import _build; print('library fresh on this box:', _build.is_fresh())"
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload cfg4 --steps 10 --warmup 3 > gpurun_out/bench_cfg4_${N}gpu.json 2> gpurun_out/bench_cfg4_${N}gpu.err
echo "cfg4 rc=$?"; tail -1 gpurun_out/bench_cfg4_${N}gpu.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'infer', round(d['infer']['value']))"
