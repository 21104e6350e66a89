#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --no-cpu --no-knn --steps 10 > gpurun_out/bench_only.json 2> gpurun_out/bench_only.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_only.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/step', round(d['ms_per_step'], 3))
print({k: round(v, 4) for k, v in d['roofline']['stage_ms_per_batch'].items()})
PY
tail -3 gpurun_out/bench_only.err
true
