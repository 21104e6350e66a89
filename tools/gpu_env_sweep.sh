#!/bin/bash
# usage: VAR=name VALS="a b c" bash tools/gpu_env_sweep.sh
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_extractor_gpu.py -x -q -m gpu --timeout 90 2>&1 | tail -2
for t in $VALS; do
  echo "== $VAR=$t"
  env $VAR=$t timeout 300 python bench.py --no-cpu --no-knn --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],4), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_batch'].items()})"
done
