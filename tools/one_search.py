"""Developer aid: a few single SearchByProjection(Cur, Last) calls on two consecutive synthetic frames (for ncu captures of
k_search_rounds / k_search)."""
import sys, numpy as np
sys.path.insert(0, '.')
from mam3slam_b200 import ORBextractor, ORBmatcher, synth
fr = synth.stream(752, 480, 0, 2)
ex = ORBextractor(1000, 1.2, 8, 20, 7)
(_, k1, d1), (_, k2, d2) = ex(fr[0]), ex(fr[1])
M = ORBmatcher(0.9, True)
F = M.frame(k2, d2, (0.0, 0.0, 752.0, 480.0))
sf = ex.GetScaleFactors()
for i in range(6):
    nm, km = M.SearchByProjection_Frame(F, np.zeros(len(k2), np.uint8), sf, np.ones(len(k1), np.uint8), k1["x"], k1["y"], k1["octave"],
                                        k1["angle"], d1, 15.0)
print("matches", nm)
