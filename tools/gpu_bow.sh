#!/bin/bash
# per-call SearchByBoW: host split + kernel durations
python tools/bow_latency.py 30 > gpurun_out/bow_lat.txt 2>&1; echo "rc $?" >> gpurun_out/bow_lat.txt
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 200 --csv --log-file gpurun_out/bow_launches.csv python tools/bow_latency.py 5 > gpurun_out/bow_ncu.log 2>&1
tail -5 gpurun_out/bow_lat.txt
grep -c k_search gpurun_out/bow_launches.csv
