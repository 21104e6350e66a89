#!/bin/bash
# native "A agents x one frame in flight" rows (tools/agents_native.cc) on the synthetic EuRoC-shaped stream
mkdir -p gpurun_out
python - <<'PY'
import numpy as np
from mam3slam_b200 import synth
fr = synth.stream(752, 480, 0, 16)
np.stack(fr).astype(np.uint8).tofile('/tmp/frames.u8')
PY
for t in 1 0; do
  timeout 600 ./mam3slam_b200/agents_native /tmp/frames.u8 752 480 16 0 ${PER:-300} ${COUNTS:-1,2,4,8,16,32} $t | tee gpurun_out/agents_native_track$t.json
done
