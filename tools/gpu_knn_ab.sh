#!/bin/bash
# same box: single-GPU k-NN with and without the fused merge epilogue, then the 2-GPU sharded search (peer stores vs NCCL)
mkdir -p gpurun_out
for f in 0 1 0 1; do
  ORBM_KNN_FUSED=$f CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-kitti --no-sparse --no-cpu > gpurun_out/knn_fused$f.json 2> gpurun_out/knn_fused$f.err
  python -c "import json; d=json.load(open('gpurun_out/knn_fused$f.json')); print('fused=$f knn ms', round(d['hamming']['knn']['ms'],4))"
done
bash tools/gpu_multi.sh 2 | grep -E "knn|N 2|rc"
