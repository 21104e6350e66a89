#!/bin/bash
# $1 = tag.  Full GPU test suite, the default bench line (with CPU baselines), the reference arm, then the ncu evidence.
TAG=${1:-r2c}
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/${TAG}_gpu_tests.log 2>&1; echo "gpu tests rc $?"; tail -3 gpurun_out/${TAG}_gpu_tests.log
timeout 1200 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"; tail -2 gpurun_out/${TAG}_bench.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc $?"
python - <<PY
import json
d=json.load(open('gpurun_out/${TAG}_bench.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/step', round(d['ms_per_step'],3), 'launches', d['gpu_launches'])
print(d['roofline'].get('stage_ms_per_batch'))
print('knn', d['hamming']['knn']['ms'], d['hamming']['knn']['roofline_int']['frac'])
L=d['latency']; print({k:v for k,v in L.items() if k not in ('search_members','agents_one_frame_in_flight','agents_one_frame_in_flight_native')}); print(json.dumps(L.get('agents_one_frame_in_flight_native'))[:1500])
print(L['agents_one_frame_in_flight']['rows'])
for r in L['search_members']['rows']: print(f"{r['member']:48s} gpu {r['gpu_call_ms']:.3f} ms  cpu {r['cpu_port_call_ms']:.3f} ms  x{r['ratio']:.1f}")
print('cpu', d.get('cpu_baseline',{}).get('value'), 'cv2', (d.get('cpu_baseline_cv2') or {}).get('value'))
r=json.load(open('gpurun_out/${TAG}_bench_ref.json')); print('ref', r.get('value'), r.get('cpu_baseline',{}).get('cores'))
PY
bash tools/gpu_ncu_r2.sh ${TAG} 2>&1 | tail -6
