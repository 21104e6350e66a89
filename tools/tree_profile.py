"""Phase split of the level-0 quadtree CTA of one frame (library built with -DORBX_TREE_PROFILE as build_variants/prof.so)."""
import ctypes as C, sys, shutil, numpy as np
sys.path.insert(0, '.')
shutil.copy('build_variants/prof.so', 'mam3slam_b200/libmam3slam_b200.so')
from mam3slam_b200 import ORBextractor, synth, _lib
L = _lib.lib()
frames = synth.stream(752, 480, 0, 8)
ex = ORBextractor(1000, 1.2, 8, 20, 7)
for i in range(5): ex(frames[i])
z = (C.c_longlong * 16)()
L.orbx_debug_tree_profile(z, 1)
N = 40
for i in range(N): ex(frames[i % 8])
L.orbx_debug_tree_profile(z, 0)
names = ["gather", "roots", "scan+proc", "split_big", "split_nodes", "apply_splits", "sorted:sort+split", "sorted:rest", "select best"]
tot = sum(z[:9])
for n, v in zip(names, z[:9]): print(f"{n:20s} {v / N / 1965:.1f} us  {100 * v / tot:.0f}%")
print("sort (warp 0)", z[9] / N / 1965, "us; split (warp 1)", z[10] / N / 1965, "us; m avg", z[11] / max(z[12], 1), "iters/frame", z[12] / N)
print("total", tot / N / 1965, "us")
