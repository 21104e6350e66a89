"""Developer aid: where a per-call SearchByBoW goes (slot rebinding vs the search call); run under ncu for the kernel durations."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from mam3slam_b200 import ORBextractor, ORBmatcher, synth
import matcher_scenarios as sc
W, H = 752, 480
ex = ORBextractor(1000, 1.2, 8, 20, 7)
f = synth.stream(W, H, 0, 2)
_, k1, d1 = ex(f[0]); _, k2, d2 = ex(f[1])
bounds = (0., 0., float(W), float(H))
fv1, fv2 = sc.feat_vec(d1), sc.feat_vec(d2)
ones1, ones2 = np.ones(len(k1), np.uint8), np.ones(len(k2), np.uint8)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
ts, tq = [], []
for it in range(reps):
    t0 = time.perf_counter()
    m = ORBmatcher(0.7, True)
    A = m.slot(0, k1, d1, bounds); B = m.slot(1, k2, d2, bounds)
    t1 = time.perf_counter()
    r = m.SearchByBoW_KF_Frame(A, fv1, ones1, B, fv2)
    t2 = time.perf_counter()
    m.close()
    ts.append(t1 - t0); tq.append(t2 - t1)
print('n1 %d n2 %d nodes %d/%d matches %d' % (len(k1), len(k2), len(fv1[0]), len(fv2[0]), int((np.asarray(r[0] if isinstance(r, tuple) else r) >= 0).sum())))
print('slot rebinding %.3f ms, search call %.3f ms (medians)' % (np.median(ts) * 1e3, np.median(tq) * 1e3))
