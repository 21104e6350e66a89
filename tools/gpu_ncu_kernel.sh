#!/bin/bash
# ncu --set full of ONE kernel (regex $KERN, skip $SKIP matching launches) of the bench command
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu ${EXTRA}"
$CMD > gpurun_out/plain3.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain3.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:${KERN} -s ${SKIP:-3} -c ${COUNT:-1} -f -o gpurun_out/${OUT:-prof_kernel} $CMD > gpurun_out/ncu_kernel.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu_kernel.log
