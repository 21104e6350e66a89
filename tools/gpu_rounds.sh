#!/bin/bash
# claim-rounds search: matcher parity + native agent rows, rounds on / off
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_matcher_gpu.py tests/test_cpp_adapter_gpu.py tests/test_agents_gpu.py -x -q -m gpu > gpurun_out/rounds_tests.log 2>&1; echo "tests rc $?"; tail -5 gpurun_out/rounds_tests.log
python - <<'PY'
import numpy as np
from mam3slam_b200 import synth
np.stack(synth.stream(752, 480, 0, 16)).astype(np.uint8).tofile('/tmp/frames.u8')
PY
for r in 1 0; do
  echo "ORBM_ROUNDS=$r"
  ORBM_ROUNDS=$r timeout 300 ./mam3slam_b200/agents_native /tmp/frames.u8 752 480 16 0 300 ${COUNTS:-1,8,16} 1
done
