#!/bin/bash
# frames in flight per GPU per step: where do the tail effects of the small-level kernels stop paying?
for b in 64 96 128 192 256; do
  echo "== batch $b"
  timeout 300 python bench.py --no-cpu --no-knn --no-kitti --steps 20 --batch $b --sets $((512 / b > 2 ? 512 / b : 3)) 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],3), {k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})"
done
