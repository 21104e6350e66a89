#!/bin/bash
python - <<'PY'
import numpy as np
from mam3slam_b200 import synth
np.stack(synth.stream(752, 480, 0, 16)).astype(np.uint8).tofile('/tmp/frames.u8')
PY
for pin in 0 1 0 1; do for t in 0 1; do ./mam3slam_b200/agents_native /tmp/frames.u8 752 480 16 0 300 1,16 $t $pin | python -c "
import json,sys
d=json.load(sys.stdin)
for r in d['rows']: print('pin', d['pinned_frames'], 'track', d['track'], r['agents'], round(r['frames_per_s']), r['median_frame_latency_ms'], r['mean_call_ms'])"; done; done
timeout 300 python -m pytest tests/test_agents_native.py -x -q 2>&1 | tail -2
