#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_extractor_gpu.py -x -q -m gpu --timeout 90 2>&1 | tail -3
for t in 64 128 256; do
  echo "== ORBX_FAST_THREADS=$t"
  ORBX_FAST_THREADS=$t timeout 300 python bench.py --no-cpu --no-knn --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), {k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})"
done
