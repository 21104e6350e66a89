#!/bin/bash
# round-2 ncu evidence.  $1 = tag (e.g. r2a).  Serial schedule, one batch of 64 frames per step; each ncu pass runs only after the
# same command has exited 0 without ncu.  (1) --set full of ONE whole batch (18 launches), (2) --set full of the k-NN kernels,
# (3) launch list (durations only) of the whole command.
TAG=${1:-r2a}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --batches-per-step 1 --no-cpu --no-kitti --no-sparse --no-latency --pipes 1 --e2e-pipes 1"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
# launches before the first timed batch: handle creation pass (14) + 8 graph-priming batches + 3 warm-up batches (18 each)
SKIP=$((14 + 11 * 18))
ncu --set full --clock-control none --import-source on -s $SKIP -c 18 -f -o gpurun_out/prof_${TAG}_step $CMD > gpurun_out/${TAG}_ncu_step.log 2>&1
echo "ncu step exit $?"
ncu --set full --clock-control none --import-source on --kernel-name "regex:k_knn" -s 6 -c 2 -f -o gpurun_out/prof_${TAG}_knn $CMD > gpurun_out/${TAG}_ncu_knn.log 2>&1
echo "ncu knn exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
echo "ncu list exit $?"; grep -c "gpu__time_duration" gpurun_out/${TAG}_launches.csv
ls -la gpurun_out/*.ncu-rep
