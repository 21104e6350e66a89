#!/bin/bash
# A/B of library variants under build_variants/: extractor parity, then the bench stage split; restores the default build
cp mam3slam_b200/libmam3slam_b200.so /tmp/default.so
for v in build_variants/*.so; do
  cp $v mam3slam_b200/libmam3slam_b200.so
  echo "== $v"
  timeout 300 python -m pytest tests/test_extractor_gpu.py -x -q -m gpu --timeout 120 2>&1 | tail -2
  timeout 300 python bench.py --no-cpu --no-knn --steps 30 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), {k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})"
done
cp /tmp/default.so mam3slam_b200/libmam3slam_b200.so
