#!/bin/bash
# round 2, first GPU pass: extractor parity with the two-pass FAST, then A/B of FAST generations / compass / occupancy
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_extractor_gpu.py tests/test_extractor_vs_ref_gpu.py tests/test_env_switches_gpu.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2a_pytest.log
cat gpurun_out/r2a_pytest.log
B="timeout 300 python bench.py --no-cpu --no-knn --no-kitti --steps 50"
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), {k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})"; }
echo "== gen1";  ORBX_FAST_GEN=1 $B 2>gpurun_out/r2a_err1.log | show
echo "== gen2";  $B 2>gpurun_out/r2a_err2.log | tee gpurun_out/r2a_bench_gen2.json | show
echo "== gen2 compass"; ORBX_FAST_COMPASS=1 $B 2>/dev/null | show
cp mam3slam_b200/libmam3slam_b200.so /tmp/default.so
cp build_variants/minb4.so mam3slam_b200/libmam3slam_b200.so
echo "== gen2 minblocks4"; $B 2>/dev/null | show
cp /tmp/default.so mam3slam_b200/libmam3slam_b200.so
tail -3 gpurun_out/r2a_err2.log
