#!/bin/bash
# FAST launch groups (cells of ordinary height launched with their own, smaller shared-memory size) vs one launch
timeout 400 python -m pytest tests/test_extractor_gpu.py -x -q -m gpu --timeout 120 2>&1 | tail -1
for gsel in 0 1 0 1; do
  echo "== ORBX_FAST_GROUPS=$gsel"
  ORBX_FAST_GROUPS=$gsel timeout 300 python bench.py --no-cpu --no-knn --steps 50 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'fast', round(d['roofline']['stage_ms_per_batch']['fast'],4), 'kitti', round(d['kitti']['extract_frames_per_s']))"
done
