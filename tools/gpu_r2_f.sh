#!/bin/bash
for x in none p a u pau; do
  STGCN_X_SKIP=$x timeout 100 python bench.py --steps 20 --warmup 5 --only-train 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('skip=$x', round(d['value']), round(d['ms_per_step'],3))"
done
