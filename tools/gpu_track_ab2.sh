#!/bin/bash
for setting in "ORBM_TRACK_ROUNDS=0" "ORBM_TRACK_ROUNDS=1 ORBM_TRACK_ROUNDS_THREADS=256" "ORBM_TRACK_ROUNDS=1 ORBM_TRACK_ROUNDS_THREADS=512" "ORBM_TRACK_ROUNDS=0" "ORBM_TRACK_ROUNDS=1 ORBM_TRACK_ROUNDS_THREADS=256" "ORBM_TRACK_ROUNDS=1 ORBM_TRACK_ROUNDS_THREADS=512"; do
  env $setting python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse --no-cpu --no-latency > /tmp/ab.json 2>/dev/null
  python -c "import json; d=json.load(open('/tmp/ab.json')); print('$setting', round(d['value']), round(d['e2e']['value']), round(d['roofline']['stage_ms_per_batch']['track'],4), round(d['roofline']['stage_ms_per_batch']['total'],4))"
done
