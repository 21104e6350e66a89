"""Developer aid: where a single-frame ORBextractor::operator() call goes (stage timing on, one frame in flight), dense and
sparse texture, plus the ncu launch list of the same calls when run under ncu."""
import sys, time, numpy as np
sys.path.insert(0, '.')
from mam3slam_b200 import ORBextractor, synth
W, H = 752, 480
ex = ORBextractor(1000, 1.2, 8, 20, 7)
for name, sparse in (("dense", False), ("sparse", True)):
    frames = [synth.frame(W, H, i, sparse=sparse) for i in range(4)]
    for f in frames: ex(f)
    t = []
    for i in range(40):
        t0 = time.perf_counter(); ex(frames[i % 4]); t.append(time.perf_counter() - t0)
    print(name, "operator() median %.3f ms  min %.3f ms" % (np.median(t) * 1e3, np.min(t) * 1e3))
    ex.set_stage_timing(True)
    acc = {}
    for i in range(10):
        ex(frames[i % 4]); ex.synchronize()
        for k, v in ex.stage_times().items(): acc[k] = acc.get(k, 0) + v / 10
    ex.set_stage_timing(False)
    print("  stages (serial, events):", {k: round(v * 1e3) for k, v in acc.items()}, "us")
