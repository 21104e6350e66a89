#!/bin/bash
for p in 2 3 4 5; do
  python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse --no-cpu --no-latency --pipes $p --e2e-pipes $p > /tmp/ab.json 2>/dev/null
  python -c "import json; d=json.load(open('/tmp/ab.json')); print('pipes $p', round(d['value']), round(d['e2e']['value']))"
done
