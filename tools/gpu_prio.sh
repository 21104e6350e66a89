#!/bin/bash
# launch priority of the latency-bound kernels (ORBX_PRIO=1) vs plain launches, for 2-4 handle pairs
timeout 300 python -m pytest tests/test_extractor_gpu.py tests/test_matcher_gpu.py -x -q -m gpu --timeout 120 2>&1 | tail -2
for prio in 0 1; do for pipes in 1 2 3 4; do
  echo "== ORBX_PRIO=$prio pipes $pipes"
  ORBX_PRIO=$prio timeout 300 python bench.py --no-cpu --no-knn --no-kitti --steps 30 --pipes $pipes --e2e-pipes 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],3))"
done; done
