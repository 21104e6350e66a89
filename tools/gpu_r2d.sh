#!/bin/bash
# round 2: new bench line (multi-batch steps, baselines, sparse leg) + env-switch parity + reference arm
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_env_switches_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc $?"; tail -3 gpurun_out/r2d_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2d_bench_ref.json 2> gpurun_out/r2d_ref.err; echo "ref rc $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2d_bench.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/step', round(d['ms_per_step'],3), 'launches', d['gpu_launches'], 'clocks', d['clocks'])
print({k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})
print('roofline', {k: d['roofline'][k] for k in ('kernel','achieved','frac','traffic','traffic_source')})
print('cpu', d.get('cpu_baseline')); print('cv2', d.get('cpu_baseline_cv2'))
print('knn', d['hamming']['knn']); print('sparse', d['sparse']); print('latency', d['latency'])
r=json.load(open('gpurun_out/r2d_bench_ref.json')); print('ref', r['value'], r['config']==d['config'])
PY
