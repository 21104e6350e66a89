#!/usr/bin/env python3
"""Micro-driver for the device-resident tracking step and the k-NN kernel (used under ncu): one extraction pass, then a few
track / knn launches.  usage: tools/bench_track.py [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from mam3slam_b200 import ORBextractor, ORBmatcher, synth
from mam3slam_b200.matcher import KnnDatabase

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B = 64
frames = synth.stream(752, 480, 0, B)
dev = torch.from_numpy(frames).cuda()
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
mt = ORBmatcher(0.9, True)
ex.extract_device(dev.data_ptr(), 752 * 480, B, 752, 480, 752)
ex.synchronize()
dk, dd, dc, stride = ex.results_batch_device()
kpm = torch.empty((B, stride), dtype=torch.int32, device="cuda")
nm = torch.empty(B, dtype=torch.int32, device="cuda")
sf = ex.GetScaleFactors()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
st = torch.cuda.current_stream().cuda_stream
for i in range(iters):
    e0.record()
    mt.track_batch_device(B, dk, dd, dc, stride, (0, 0, 752, 480), sf, 15.0, kpm.data_ptr(), nm.data_ptr(), stream=st)
    e1.record()
    torch.cuda.synchronize()
    print("track ms", e0.elapsed_time(e1), "matches/frame", float(nm[1:].float().mean()), "evals", mt.track_evaluations())
g = torch.Generator(device="cuda").manual_seed(1)
db = torch.randint(0, 256, (1_000_000, 32), dtype=torch.uint8, device="cuda", generator=g)
q = db[:2000].clone()
D = KnnDatabase(mt, None, device_ptr=db.data_ptr(), n=1_000_000)
od = torch.empty((2000, 2), dtype=torch.int16, device="cuda")
oi = torch.empty((2000, 2), dtype=torch.int32, device="cuda")
for i in range(iters):
    e0.record()
    mt.knn_device(D, q.data_ptr(), 2000, 2, od.data_ptr(), oi.data_ptr(), st)
    e1.record()
    torch.cuda.synchronize()
    print("knn ms", e0.elapsed_time(e1))
