#!/bin/bash
timeout 300 python -m pytest tests/test_matcher_gpu.py -x -q -m gpu --timeout 120 -k "knn" 2>&1 | tail -2
for q in 1 2; do
  echo "== ORBM_KNN_QPT=$q"
  ORBM_KNN_QPT=$q timeout 300 python bench.py --no-cpu --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['hamming']['knn']; print('knn ms', round(k['ms'],3), 'G evals/s', round(k['evaluations_per_s']/1e9,1), k['planted_neighbours_found'])"
done
