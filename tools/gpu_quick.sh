#!/bin/bash
# extractor parity + a short bench (stage times) after a kernel change
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_extractor_gpu.py tests/test_extractor_vs_ref_gpu.py tests/test_env_switches_gpu.py -x -q -m gpu > gpurun_out/quick_tests.log 2>&1; echo "tests rc $?"; tail -3 gpurun_out/quick_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse --no-cpu > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err; echo "bench rc $?"; tail -3 gpurun_out/quick_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/quick_bench.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/step', round(d['ms_per_step'],3))
print(d['roofline'].get('stage_ms_per_batch'))
print({k:v for k,v in d['latency'].items() if k!='search_members'})
PY
