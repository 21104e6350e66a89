"""Developer aid: one EuRoC-shaped frame through the extractor, compared with the oracle (run on the GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from mam3slam_b200 import ORBextractor, synth
img = synth.frame(752, 480, 0)
ex = ORBextractor(1000, 1.2, 8, 20, 7)
mono, kps, desc = ex(img)
orc = oracle.Extractor(1000, 1.2, 8, 20, 7)
omono, okps, odesc = orc(img)
for l in range(8):
    oc = orc.level_points(l)[:, :3].astype(np.int32)
    c = ex.candidates(l)
    print("level", l, "cand", len(c), len(oc), "equal" if np.array_equal(c, oc) else "DIFFER")
print("ok", mono, omono, len(kps), len(okps), np.array_equal(desc, odesc) if len(kps) == len(okps) else None)
