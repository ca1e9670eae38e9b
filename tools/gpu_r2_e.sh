#!/bin/bash
mkdir -p gpurun_out
for ch in default 2 4 8; do
  if [ "$ch" = default ]; then unset NCCL_MAX_NCHANNELS; else export NCCL_MAX_NCHANNELS=$ch; fi
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 295$((RANDOM%90+10)) bench.py --gpus 2 --steps 20 --warmup 5 --only-train 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('NCCL_MAX_NCHANNELS=$ch', round(d['value']), round(d['ms_per_step'],3))"
done
timeout 100 python bench.py --steps 20 --warmup 5 --only-train 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('1 GPU', round(d['value']), round(d['ms_per_step'],3))"
