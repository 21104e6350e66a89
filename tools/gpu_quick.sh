#!/bin/bash
# quick GPU check: extractor + matcher parity, bench line
mkdir -p gpurun_out
timeout 400 python -m pytest tests -x -q -m gpu --timeout 90 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_quick.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/step', round(d['ms_per_step'], 3))
print({k: round(v, 4) for k, v in d['roofline']['stage_ms_per_batch'].items()})
print('knn ms', d.get('hamming', {}).get('knn', {}).get('ms'))
print('latency', d.get('latency'))
print('kitti', {k: (round(v, 3) if isinstance(v, float) else v) for k, v in (d.get('kitti') or {}).items() if k != 'workload'})
PY
tail -3 gpurun_out/bench_quick.err
[ -x tools/microbench ] && [ -n "$MICRO" ] && tools/microbench > gpurun_out/microbench.txt && tail -8 gpurun_out/microbench.txt
true
