import sys; sys.path.insert(0,'.')
import numpy as np
from mam3slam_b200 import ORBextractor, synth
import oracle
img = synth.frame(752,480,0)
ex = ORBextractor(1000,1.2,8,20,7); orc = oracle.Extractor(1000,1.2,8,20,7)
mono,kps,desc = ex(img); omono, okps, odesc = orc(img)
for l in range(8):
    p = np.array_equal(ex.image_pyramid_level(l), orc.level_pixels(l))
    c = ex.candidates(l); oc = orc.level_points(l)[:, :3].astype(np.int32)
    b = np.array_equal(ex.blurred_level(l), orc.level_pixels(l,'blurred'))
    k = ex.level_keypoints(l); ok = orc.level_points(l, True)[:, :3].astype(np.int32); ok[:, :2]-=16
    sc = set(map(tuple,c.tolist())); so=set(map(tuple,oc.tolist()))
    print(l, 'pyr',p,'blur',b,'cand',len(c),len(oc),len(sc&so), 'kept', len(k), len(ok), np.array_equal(k,ok))
    if l==0:
        print(c[:10].tolist()); print(oc[:10].tolist())
        sxy = {(a,b_):s for a,b_,s in c.tolist()}; 
        diffs=[(a,b_,s,sxy.get((a,b_))) for a,b_,s in oc.tolist() if sxy.get((a,b_))!=s][:10]; print('diffs',diffs)
print(mono,omono,len(kps),len(okps))
