"""Developer aid: resolver / producer cycle counters of the tracking kernel (library built with EXTRA=-DORBM_STATS)."""
import ctypes, sys, numpy as np, torch
sys.path.insert(0, '.')
from mam3slam_b200 import ORBextractor, ORBmatcher, _lib, synth
import os
B, W, H = int(os.environ.get('B', 64)), 752, 480
frames = np.stack(synth.stream(W, H, 0, B))
dev = torch.from_numpy(frames).cuda()
ex = ORBextractor(1000, 1.2, 8, 20, 7, device=0, max_batch=B)
mt = ORBmatcher(0.9, True, device=0)
sf = ex.GetScaleFactors()
ex.extract_device(dev.data_ptr(), W * H, B, W, H, W, (0, 1000), 0)
torch.cuda.synchronize()
dk, dd, dc, stride = ex.results_batch_device()
kpm = torch.empty((B, stride), dtype=torch.int32, device="cuda")
nmt = torch.empty(B, dtype=torch.int32, device="cuda")
for it in range(3):
    mt.track_batch_device(B, dk, dd, dc, stride, (0., 0., float(W), float(H)), sf, 15.0, kpm.data_ptr(), nmt.data_ptr())
    torch.cuda.synchronize()
c = (ctypes.c_ulonglong * 16)()
L = mt._L
L.orbm_debug_track_counters.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
print('rc', L.orbm_debug_track_counters(mt._h, c))
v = list(c)
n = max(v[6], 1)
print('evals', v[0], 'CTAs', v[6], 'matches', int(nmt.sum()))
print('per CTA: prologue %.0f, main loop (tid 0) %.0f cycles' % (v[10]/n, v[9]/n))
print('resolver per CTA: wait %.0f resolve %.0f cycles; iterations %.1f re-walks %.1f' % (v[1]/n, v[3]/n, v[4]/n, v[5]/n))
print('producers per warp: wait %.0f work %.0f cycles' % (v[7]/n/7, v[8]/n/7))
print('resolver phases per CTA: pick %.0f conflict %.0f commit %.0f' % (v[11]/n, v[12]/n, v[13]/n))
