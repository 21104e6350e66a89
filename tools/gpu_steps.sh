#!/bin/bash
for st in 20 50 200 1000 200; do
  echo "== steps $st"
  timeout 300 python bench.py --no-cpu --no-knn --no-kitti --steps $st 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],4), d['clocks']['samples'])"
done
