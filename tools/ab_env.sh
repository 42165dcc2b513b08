#!/bin/bash
# Same-box A/B of an environment switch: `gpurun -- 'bash tools/ab_env.sh B200DP_L2_PERSIST 0 1 [bench.py args]'`
var=$1; a=$2; b=$3; shift 3
for round in 1 2; do
  for v in $a $b; do
    echo -n "$var=$v round$round "
    env $var=$v timeout 100 python bench.py --steps 1500 --warmup 30 --skip-e2e "$@" 2>&1 | tail -1 | python -c "
import sys,json
l=sys.stdin.read()
try:
    d=json.loads(l); print(d['value'], d['ms_per_step'])
except Exception: print('FAILED', l[-300:])"
  done
done
