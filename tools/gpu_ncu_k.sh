#!/bin/bash
# ncu --set full of the kernels matching $1 (regex), $2 launches, from the serial bench command; then the launch list
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-knn --no-kitti --pipes 1 --e2e-pipes 1"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on --kernel-name "regex:$1" -s ${SKIP:-8} -c ${2:-4} -f -o gpurun_out/${OUT:-prof_k} $CMD > gpurun_out/ncu_k.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_k.log; ls -la gpurun_out/*.ncu-rep
