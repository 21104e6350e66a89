#!/bin/bash
# k-NN paths after a change: parity tests + one timing of the 1 M x 2000 search
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_matcher_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu -k "knn" > gpurun_out/knn_tests.log 2>&1; echo "tests rc $?"; tail -4 gpurun_out/knn_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-kitti --no-sparse --no-cpu > gpurun_out/knn_bench.json 2> gpurun_out/knn_bench.err; echo "bench rc $?"; tail -3 gpurun_out/knn_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/knn_bench.json'))
k=d['hamming']['knn']; print('knn ms', k['ms'], 'frac', k['roofline_int']['frac'])
print('e2e', d['e2e'])
PY
