#!/bin/bash
# ncu --set full capture of one whole step (16 launches after 3 warm-up steps) + the k-NN kernels
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --pipes 1 --e2e-pipes 1"
$CMD > gpurun_out/plain2.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain2.log; exit 1; }
ncu --set full --clock-control none --import-source on -s ${SKIP:-48} -c ${COUNT:-16} -f -o gpurun_out/${OUT:-prof_r1b} $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep
