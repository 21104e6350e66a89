"""Single-frame latency of the drop-in calls (one frame at a time, like Tracking does): operator() + SearchByProjection."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
from mam3slam_b200 import ORBextractor, ORBmatcher, synth
W, H = 752, 480
frames = synth.stream(W, H, 0, 8)
pin = [torch.from_numpy(f.copy()).pin_memory().numpy() for f in frames]
ex = ORBextractor(1000, 1.2, 8, 20, 7)
m = ORBmatcher(0.9, True)
sf = ex.GetScaleFactors()
for name, src in (("pageable", frames), ("pinned", pin)):
    for i in range(10): ex(src[i % 8])
    t0 = time.perf_counter(); N = 200
    for i in range(N): mono, kps, desc = ex(src[i % 8])
    dt = (time.perf_counter() - t0) / N
    print(f"operator() {name}: {dt*1e3:.3f} ms/frame  ({1/dt:.0f} frames/s)")
mono, k0, d0 = ex(frames[0]); mono, k1, d1 = ex(frames[1])
F = m.frame(k1, d1, (0, 0, W, H))
occ = np.zeros(len(k1), np.uint8); valid = np.ones(len(k0), np.uint8)
for i in range(10): m.SearchByProjection_Frame(F, occ, sf, valid, k0["x"], k0["y"], k0["octave"], k0["angle"], d0, 15.0)
t0 = time.perf_counter(); N = 200
for i in range(N):
    F.update(k1, d1, (0, 0, W, H)) if hasattr(F, "update") else None
    nm, km = m.SearchByProjection_Frame(F, occ, sf, valid, k0["x"], k0["y"], k0["octave"], k0["angle"], d0, 15.0)
dt = (time.perf_counter() - t0) / N
print(f"frame update + SearchByProjection(Cur, Last): {dt*1e3:.3f} ms  matches {nm}")
