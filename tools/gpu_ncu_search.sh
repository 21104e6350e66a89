#!/bin/bash
# ncu --set full of the single-call tracking search: claim rounds (default) and the in-order kernel (ORBM_ROUNDS=0)
TAG=${1:-r2g}
mkdir -p gpurun_out
python tools/one_search.py || exit 1
ncu --set full --clock-control none --import-source on --kernel-name "regex:k_search_rounds" -s 3 -c 2 -f -o gpurun_out/prof_${TAG}_search_rounds python tools/one_search.py > gpurun_out/${TAG}_ncu_search.log 2>&1; echo "ncu rounds exit $?"
ORBM_ROUNDS=0 ncu --set full --clock-control none --import-source on --kernel-name "regex:k_search" -s 3 -c 2 -f -o gpurun_out/prof_${TAG}_search_inorder python tools/one_search.py >> gpurun_out/${TAG}_ncu_search.log 2>&1; echo "ncu in-order exit $?"
