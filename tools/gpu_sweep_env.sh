#!/bin/bash
# value / e2e / stage split of the bench under a list of env settings: tools/gpu_sweep_env.sh "A=1" "A=2 B=0" ...
B="timeout 300 python bench.py --no-cpu --no-knn --no-kitti --no-sparse --steps 10"
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/batch', round(d['ms_per_step']/d['config']['batches_per_step'],4), {k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})"; }
echo "== default"; $B 2>/dev/null | show
for e in "$@"; do echo "== $e"; env $e $B 2>/dev/null | show; done
