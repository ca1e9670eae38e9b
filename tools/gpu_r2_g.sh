#!/bin/bash
# N-GPU scaling points: tools/gpu_r2_g.sh N
N=${1:-8}; mkdir -p gpurun_out
for wl in cfg4 cfg5; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$((RANDOM%9)) bench.py --gpus $N --workload $wl --steps 10 --warmup 3 > gpurun_out/bench_${wl}_${N}gpu.json 2> gpurun_out/bench_${wl}_${N}gpu.err
  echo "$wl rc=$?"; tail -1 gpurun_out/bench_${wl}_${N}gpu.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'infer', round(d['infer']['value']))"
done
