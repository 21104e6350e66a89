#!/bin/bash
# A/B of one environment switch of the library: VAR=name VALS="0 1" tools/gpu_env_ab.sh
for rep in 1 2; do for v in $VALS; do
  echo "== $VAR=$v"
  env $VAR=$v timeout 300 python bench.py --no-cpu --no-knn --no-kitti --steps 100 $EXTRA 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],4), 'latency', round(d['latency']['operator_call_ms'],4))"
done; done
