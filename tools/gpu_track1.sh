#!/bin/bash
# single-pair tracking: native per-call times + resolver / producer counters of one CTA (stats build)
mkdir -p gpurun_out
python - <<'PY'
import numpy as np
from mam3slam_b200 import synth
np.stack(synth.stream(752, 480, 0, 16)).astype(np.uint8).tofile('/tmp/frames.u8')
PY
./mam3slam_b200/agents_native /tmp/frames.u8 752 480 16 0 300 1 1
cp mam3slam_b200/libmam3slam_b200.so /tmp/base.so
cp build_variants/stats.so mam3slam_b200/libmam3slam_b200.so
B=2 python tools/track_stats.py
B=64 python tools/track_stats.py
cp /tmp/base.so mam3slam_b200/libmam3slam_b200.so
