#!/bin/bash
# same-box A/B of an environment switch: tools/gpu_ab.sh VAR [rounds]; alternates runs without / with VAR=1
VAR=${1:-STGCN_NO_TRAIN_PREP}; N=${2:-3}
for i in $(seq $N); do
  for v in 0 1; do
    if [ $v = 1 ]; then export $VAR=1; else unset $VAR; fi
    timeout 100 python bench.py --steps 30 --warmup 5 --only-train 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$VAR=$v', round(d['ms_per_step'],3), d['gpu_launches']//30)"
  done
done
