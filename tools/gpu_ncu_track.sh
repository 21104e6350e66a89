#!/bin/bash
mkdir -p gpurun_out
python tools/bench_track.py 3 > gpurun_out/track_plain.log 2>&1 && cat gpurun_out/track_plain.log &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_track|k_knn" --csv --log-file gpurun_out/track_launches.csv python tools/bench_track.py 2 > gpurun_out/track_ncu.log 2>&1
grep -E "k_track|k_knn" gpurun_out/track_launches.csv | awk -F'","' '{print $5, $NF}' | head -30
