#!/bin/bash
timeout 600 python -m pytest tests/test_env_switches_gpu.py::test_matcher_resolver_switches_do_not_change_results -x -q -m gpu 2>&1 | tail -2
for r in 1 0 1 0; do
  ORBM_TRACK_ROUNDS=$r python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse --no-cpu --no-latency > /tmp/ab.json 2>/dev/null
  python -c "import json; d=json.load(open('/tmp/ab.json')); print('ORBM_TRACK_ROUNDS=$r', round(d['value']), round(d['e2e']['value']), round(d['roofline']['stage_ms_per_batch']['track'],4), round(d['roofline']['stage_ms_per_batch']['total'],4))"
done
