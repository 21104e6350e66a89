"""GPU parity of the extractor against the reference's OWN source (oracle/_ref/liborb_ref.so = src/ORBextractor.cc
compiled unmodified against the header shim; prebuilt in the container, it travels to the GPU box with the snapshot).

The oracle restatement is out of the loop here: CUDA result == reference-class result, bit for bit, through the C ABI.
"""
import numpy as np
import pytest

from mam3slam_b200 import ORBextractor, synth
from oracle import ref as oref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oref.available(), reason="oracle/_ref not built")]


def _check(img, nfeatures, lapping=(0, 1000), levels=8, scale=1.2, ini=20, mn=7):
    ex = ORBextractor(nfeatures, scale, levels, ini, mn)
    ref = oref.Extractor(nfeatures, scale, levels, ini, mn)
    mono, kps, desc = ex(img, None, lapping)
    rmono, rkps, rdesc = ref(img, lapping)
    assert mono == rmono and len(kps) == len(rkps)
    for name, col in (("x", 0), ("y", 1), ("size", 2), ("angle", 3), ("response", 4)):
        assert np.array_equal(kps[name].view(np.uint32), rkps[:, col].view(np.uint32)), name      # bit patterns
    assert np.array_equal(kps["octave"], rkps[:, 5].astype(np.int32))
    assert np.array_equal(desc, rdesc)
    for l in range(levels):
        assert np.array_equal(ex.image_pyramid_level(l, padded=True), ref.level_pixels(l, "padded")), f"mvImagePyramid[{l}]"
        rk = ref.level_keypoints(img if l == 0 else None, l)[:, :3].astype(np.int32)
        rk[:, :2] -= 16
        assert np.array_equal(ex.level_keypoints(l), rk), f"DistributeOctTree output of level {l}"
    ex.close()
    return len(kps)


def test_baseline_configs():
    assert _check(synth.frame(752, 480, 0), 1000) >= 1000                 # config 1/2: EuRoC mono
    assert _check(synth.frame(1241, 376, 1), 2000) >= 2000                # config 3: KITTI shape, lapping split at x = 1000
    _check(synth.frame(752, 480, 2), 5000)                                # the 5x initialisation extractor (src/Tracking.cc:606)


def test_streams_and_shapes():
    for img in synth.stream(752, 480, 21, 3):
        _check(img, 1000)
    _check(synth.frame(333, 251, 3), 500)
    _check(synth.frame(640, 480, 4), 1200, levels=3, scale=2.0)           # exact-2x levels (INTER_AREA reroute)
    _check(synth.frame(330, 440, 5), 700, levels=6)                       # portrait: one quadtree root
    _check(synth.frame(752, 480, 6), 1000, lapping=(200, 500))
    _check(synth.frame(512, 384, 8), 800, levels=10, scale=1.1, ini=15, mn=5)
    _check(synth.frame(501, 335, 13), 400, levels=5)                      # unaligned rows


def test_degenerate_textures():
    rng = np.random.default_rng(11)
    w, h = 420, 310
    _check(rng.integers(0, 256, (h, w), dtype=np.uint8), 600)
    _check((120 + (synth.frame(w, h, 3).astype(np.int32) - 128) // 12).astype(np.uint8), 600)
    yy, xx = np.mgrid[0:h, 0:w]
    _check((((xx // 9 + yy // 9) & 1) * 200 + 20).astype(np.uint8), 600)
    sparse = np.full((h, w), 30, np.uint8)
    for _ in range(25):
        x, y = int(rng.integers(30, w - 40)), int(rng.integers(30, h - 40))
        sparse[y:y + 9, x:x + 9] = 220
    _check(sparse, 600)
