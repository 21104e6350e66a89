#!/bin/bash
python tools/one_frame.py 2>&1 | tail -1
timeout 1200 python -m pytest tests/test_extractor_gpu.py tests/test_extractor_vs_ref_gpu.py tests/test_env_switches_gpu.py -x -q -m gpu 2>&1 | tail -4
bash tools/gpu_sweep_env.sh "ORBX_PYR_TMA=0"
