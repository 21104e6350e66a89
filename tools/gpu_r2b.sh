#!/bin/bash
# parity of the extractor with the three-pass FAST, then gen1 / gen2 bench stage split
mkdir -p gpurun_out
python tools/one_frame.py 2>&1 | tail -10
ORBX_FAST_TMA=0 python tools/one_frame.py 2>&1 | tail -1
timeout 900 python -m pytest tests/test_extractor_gpu.py tests/test_extractor_vs_ref_gpu.py tests/test_env_switches_gpu.py -x -q -m gpu 2>&1 | tail -5
B="timeout 300 python bench.py --no-cpu --no-knn --no-kitti --steps 50"
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), {k: round(v,4) for k,v in d['roofline']['stage_ms_per_batch'].items()})"; }
echo "== gen1";  ORBX_FAST_GEN=1 $B 2>/dev/null | show
echo "== gen2";  $B 2>gpurun_out/r2b_err2.log | show
tail -3 gpurun_out/r2b_err2.log
