#!/bin/bash
# sweep of the schedule knobs of bench.py (handle pairs on the device leg, host threads on the e2e leg, batch size)
mkdir -p gpurun_out
for cfg in "--pipes 2 --e2e-pipes 3" "--pipes 3 --e2e-pipes 3" "--pipes 4 --e2e-pipes 4" "--pipes 3 --e2e-pipes 6" "--pipes 2 --e2e-pipes 3 --batch 128" "--pipes 3 --e2e-pipes 4 --batch 32"; do
  echo "== $cfg"
  timeout 300 python bench.py --no-cpu --no-knn --steps 30 $cfg 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],3))"
done
nproc
