#!/bin/bash
# N-GPU bench under torchrun ($1 = N): sharded k-NN over peer stores vs the NCCL variant, e2e with its H2D ceiling
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-kitti --no-sparse --no-cpu > gpurun_out/multi_n$N.json 2> gpurun_out/multi_n$N.err; echo "bench rc $?"; tail -5 gpurun_out/multi_n$N.err
python - <<PY
import json
d=json.load(open('gpurun_out/multi_n$N.json'))
k=d['hamming']['knn']
print('N', d['n_gpus'], 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ceiling', round(d['e2e']['ceiling_frames_per_s']), 'GB/s', round(d['e2e']['ceiling_h2d_gbs'],1))
print('knn peer ms', round(k['ms'],4), 'nccl ms', k['ms_nccl_allgather_variant'], 'found', k['planted_neighbours_found'])
print(d['e2e']['host_placement'])
PY
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1; head -12 gpurun_out/topo_n$N.txt
