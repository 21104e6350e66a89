#!/bin/bash
python - <<'PY'
import numpy as np
from mam3slam_b200 import synth
np.stack(synth.stream(752, 480, 0, 16)).astype(np.uint8).tofile('/tmp/frames.u8')
PY
cp mam3slam_b200/libmam3slam_b200.so /tmp/base.so
cp build_variants/stats.so mam3slam_b200/libmam3slam_b200.so
./mam3slam_b200/agents_native /tmp/frames.u8 752 480 16 0 6 1 1 2>&1 | head -14
cp /tmp/base.so mam3slam_b200/libmam3slam_b200.so
