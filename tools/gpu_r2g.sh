#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -4
timeout 900 python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc $?"; tail -3 gpurun_out/r2g_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2g_bench.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/step', round(d['ms_per_step'],3))
L=d['latency']; print({k:v for k,v in L.items() if k!='search_members'})
for r in L['search_members']['rows']: print(f"{r['member']:48s} gpu {r['gpu_call_ms']:.3f} ms  cpu {r['cpu_port_call_ms']:.3f} ms  x{r['ratio']:.1f}")
PY
