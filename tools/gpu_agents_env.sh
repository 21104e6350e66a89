#!/bin/bash
# native agent rows under environment switches: each argument is one "VAR=value[,VAR=value]" setting
python - <<'PY'
import numpy as np
from mam3slam_b200 import synth
np.stack(synth.stream(752, 480, 0, 16)).astype(np.uint8).tofile('/tmp/frames.u8')
PY
for rep in 1 2; do
for setting in "ORBX_NOOP=1" "$@"; do
  echo "== $setting"
  env $(echo $setting | tr ',' ' ') ./mam3slam_b200/agents_native /tmp/frames.u8 752 480 16 0 300 ${COUNTS:-1,16} ${TRACK:-0} | python -c "
import json,sys
d=json.load(sys.stdin)
for r in d['rows']: print(r['agents'], round(r['frames_per_s']), r['median_frame_latency_ms'], r['mean_call_ms'])"
done; done
