"""Developer aid: the Python latency leg of bench.py alone (operator() and SearchByProjection per call, pinned and pageable image)."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
sys.argv = ['bench.py']
import bench
frames = bench.make_frames(0, 16) if hasattr(bench, 'make_frames') else None
from mam3slam_b200 import synth, ORBextractor
fr = np.stack(synth.stream(752, 480, 0, 16))
print(bench.latency_bench(None, torch, 0, fr))
ex = ORBextractor(1000, 1.2, 8, 20, 7)
for name, arr in (("pinned", torch.from_numpy(fr[:8].copy()).pin_memory().numpy()), ("pageable", fr[:8].copy())):
    t = []
    for i in range(210):
        t0 = time.perf_counter(); ex(arr[i % 8]); t.append(time.perf_counter() - t0)
    print(name, "operator() median %.4f ms" % (np.median(t[10:]) * 1e3))
