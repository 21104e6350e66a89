"""Per-stage GPU time of ONE frame (the latency regime a single agent's Tracking thread sees) + host-side split of the call."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
from mam3slam_b200 import ORBextractor, ORBmatcher, synth
W, H = 752, 480
frames = synth.stream(W, H, 0, 8)
pin = torch.from_numpy(frames.copy()).pin_memory().numpy()
ex = ORBextractor(1000, 1.2, 8, 20, 7)
for i in range(10): ex(pin[i % 8])
ex.set_stage_timing(True)
acc = None
for i in range(50):
    ex(pin[i % 8])
    st = ex.stage_times()
    acc = st if acc is None else {k: acc[k] + v for k, v in st.items()}
print("stage ms per frame:", {k: round(v / 50, 4) for k, v in acc.items()})
ex.set_stage_timing(False)
for i in range(10): ex(pin[i % 8])
t0 = time.perf_counter()
for i in range(200): ex(pin[i % 8])
print("operator() ms:", (time.perf_counter() - t0) / 200 * 1e3)
m = ORBmatcher(0.9, True)
sf = ex.GetScaleFactors()
_, k0, d0 = ex(pin[0]); _, k1, d1 = ex(pin[1])
F = m.frame(k1, d1, (0, 0, W, H))
occ = np.zeros(len(k1), np.uint8); valid = np.ones(len(k0), np.uint8)
x, y, o, a = k0["x"].copy(), k0["y"].copy(), k0["octave"].copy(), k0["angle"].copy()
for name, upd in (("update+search", True), ("search only", False)):
    for i in range(10): m.SearchByProjection_Frame(F, occ, sf, valid, x, y, o, a, d0, 15.0)
    t0 = time.perf_counter()
    for i in range(200):
        if upd: F.update(k1, d1, (0, 0, W, H))
        m.SearchByProjection_Frame(F, occ, sf, valid, x, y, o, a, d0, 15.0)
    print(name, "ms:", (time.perf_counter() - t0) / 200 * 1e3)
