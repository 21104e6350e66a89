#!/bin/bash
for ov in 0 1; do for p in 1 2; do
  echo "== overlap $ov pipes $p"
  ORBX_OVERLAP=$ov timeout 300 python bench.py --no-cpu --no-knn --steps 20 --pipes $p 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],4))"
done; done
