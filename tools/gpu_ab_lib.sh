#!/bin/bash
# A/B of two library builds under one bench on one box: $1 = variant .so (build_variants/...), rest = bench args
V=$1; shift
cp mam3slam_b200/libmam3slam_b200.so /tmp/base.so
for rep in 1 2; do
  for lib in /tmp/base.so $V; do
    cp $lib mam3slam_b200/libmam3slam_b200.so
    python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse --no-cpu --no-latency "$@" > /tmp/ab.json 2>/dev/null
    python -c "import json; d=json.load(open('/tmp/ab.json')); print('$lib', round(d['value']), round(d['e2e']['value']), round(d['roofline']['stage_ms_per_batch']['track'],4), round(d['roofline']['stage_ms_per_batch']['octree'],4))"
  done
done
cp /tmp/base.so mam3slam_b200/libmam3slam_b200.so
