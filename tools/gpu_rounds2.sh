#!/bin/bash
# claim rounds: per-round counters (stats build), parity of everything that tracks, bench A/B rounds on / off
mkdir -p gpurun_out
bash tools/gpu_rounds_dbg.sh 2>&1 | head -8
timeout 1500 python -m pytest tests/test_matcher_gpu.py tests/test_cpp_adapter_gpu.py tests/test_agents_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/rounds_tests.log 2>&1; echo "tests rc $?"; tail -5 gpurun_out/rounds_tests.log
for r in 1 0 1 0; do
  ORBM_ROUNDS=$r python bench.py --steps 10 --warmup 3 --no-kitti --no-knn --no-sparse --no-cpu --no-latency > /tmp/ab.json 2>/dev/null
  python -c "import json; d=json.load(open('/tmp/ab.json')); print('ORBM_ROUNDS=$r', round(d['value']), round(d['e2e']['value']), d['roofline']['stage_ms_per_batch'])"
done
ORBM_ROUNDS=1 timeout 300 ./mam3slam_b200/agents_native /tmp/frames.u8 752 480 16 0 300 1,16 1
