#!/bin/bash
# A/B of an environment switch under one bench on one box: $1 = "VAR=value"
for rep in 1 2; do
  for e in "ORBX_NOOP=1" "$1"; do
    env $e python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse --no-cpu --no-latency > /tmp/ab.json 2>/dev/null
    python -c "import json; d=json.load(open('/tmp/ab.json')); s=d['roofline']['stage_ms_per_batch']; print('$e', round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in s.items()})"
  done
done
