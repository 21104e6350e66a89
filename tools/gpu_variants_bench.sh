#!/bin/bash
for v in build_variants/*.so; do
  cp $v mam3slam_b200/libmam3slam_b200.so
  echo "== $v"; timeout 300 python bench.py --no-cpu --no-knn --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), {k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})"
done
