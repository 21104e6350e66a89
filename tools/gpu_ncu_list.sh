#!/bin/bash
# ncu launch list (per-launch durations) of the bench command; run only after the plain command exits 0
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --pipes 1 --e2e-pipes 1"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu exit $?"; grep -c "gpu__time_duration" gpurun_out/launches.csv
