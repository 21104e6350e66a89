"""CPU: the oracle restatement (oracle/cpp) against the reference's OWN extractor source.

oracle/_ref/liborb_ref.so is /root/reference/src/ORBextractor.cc compiled UNMODIFIED against the header shim
oracle/ref_shim/ (recipe: oracle/Makefile `ref`).  The constructor tables, the FAST grid loop, DistributeOctTree with
its std::list / std::sort(compareNodes), IC_Angle, computeOrbDescriptor and operator()'s lapping assembly are the
reference's own code; only cv::Mat and the five pixel primitives come from the shim, and those primitives are the
cv2-pinned ones (tests/test_oracle_golden.py, tests/test_oracle_vs_cv2.py).  Everything must agree EXACTLY: keypoints,
order, responses, octaves, sizes, angles, descriptors, monoIndex, every pyramid level incl. its 19-px ring.
"""
import numpy as np
import pytest

from mam3slam_b200 import synth
from oracle import ref as oref

pytestmark = pytest.mark.skipif(not oref.available(), reason="oracle/_ref not built and /root/reference absent")


def _same(a, b, what=""):
    mono_a, k_a, d_a = a
    mono_b, k_b, d_b = b
    assert mono_a == mono_b, what
    assert k_a.shape == k_b.shape, (what, k_a.shape, k_b.shape)
    assert np.array_equal(k_a.view(np.uint32), k_b.view(np.uint32)), what      # bit patterns, incl. angles
    assert np.array_equal(d_a, d_b), what


CASES = [
    # (name, w, h, seed, nfeatures, scale, levels, ini, min, lapping)
    ("euroc", 752, 480, 0, 1000, 1.2, 8, 20, 7, (0, 1000)),
    ("euroc-b", 752, 480, 7, 1000, 1.2, 8, 20, 7, (0, 1000)),
    ("kitti-lapping", 1241, 376, 1, 2000, 1.2, 8, 20, 7, (0, 1000)),
    ("init-5x", 752, 480, 2, 5000, 1.2, 8, 20, 7, (0, 1000)),
    ("odd-shape", 333, 251, 3, 500, 1.2, 8, 20, 7, (0, 1000)),
    ("exact-2x", 640, 480, 4, 1200, 2.0, 3, 20, 7, (0, 1000)),
    ("portrait", 330, 440, 5, 700, 1.2, 6, 20, 7, (0, 1000)),     # nIni = round(w/h) = 1; below 0.5 the reference itself indexes an empty root vector (:559-588)
    ("lap-window", 752, 480, 6, 1000, 1.2, 8, 20, 7, (200, 500)),
    ("scale-1.1", 512, 384, 8, 800, 1.1, 10, 15, 5, (0, 1000)),
    ("one-level", 400, 300, 9, 300, 1.2, 1, 20, 7, (0, 1000)),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_restatement_equals_reference_source(oracle_mod, case):
    _, w, h, seed, nf, sc, lv, ini, mn, lap = case
    img = synth.frame(w, h, seed)
    orc = oracle_mod.Extractor(nf, sc, lv, ini, mn)
    ref = oref.Extractor(nf, sc, lv, ini, mn)
    ta, tb = orc.tables(), ref.tables()
    for k in ta:
        assert np.array_equal(ta[k], tb[k]), k
    a, b = orc(img, lap), ref(img, lap)
    assert len(a[1]) > 0
    _same(a, b, case[0])
    for l in range(lv):
        assert orc.level_size(l) == ref.level_size(l)
        assert np.array_equal(orc.level_pixels(l), ref.level_pixels(l)), f"level {l}"
        assert np.array_equal(orc.level_pixels(l, "padded"), ref.level_pixels(l, "padded")), f"padded level {l}"
    # per-level kept lists in the reference's list order, before rescaling: x, y, response, angle
    for l in range(lv):
        rk = ref.level_keypoints(img if l == 0 else None, l)
        ok = orc.level_points(l, kept=True)
        assert np.array_equal(rk.view(np.uint32), ok.view(np.uint32)), f"kept list of level {l}"


@pytest.mark.parametrize("kind", ["flat", "noise", "low-contrast", "checker", "sparse"])
def test_degenerate_inputs(oracle_mod, kind):
    rng = np.random.default_rng(11)
    w, h = 420, 310
    if kind == "flat":
        img = np.full((h, w), 77, np.uint8)
    elif kind == "noise":
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    elif kind == "low-contrast":
        img = (120 + (synth.frame(w, h, 3).astype(np.int32) - 128) // 12).astype(np.uint8)      # only minThFAST fires
    elif kind == "checker":
        yy, xx = np.mgrid[0:h, 0:w]
        img = (((xx // 9 + yy // 9) & 1) * 200 + 20).astype(np.uint8)                            # ties everywhere
    else:
        img = np.full((h, w), 30, np.uint8)
        for _ in range(25):
            x, y = int(rng.integers(30, w - 40)), int(rng.integers(30, h - 40))
            img[y:y + 9, x:x + 9] = 220
    a = oracle_mod.Extractor(600, 1.2, 8, 20, 7)(img)
    b = oref.Extractor(600, 1.2, 8, 20, 7)(img)
    _same(a, b, kind)
    if kind == "flat":
        assert len(a[1]) == 0 and a[0] == 0


def test_empty_image_returns_minus_one(oracle_mod):
    # src/ORBextractor.cc:1090-1091
    ref = oref.Extractor(100, 1.2, 4, 20, 7)
    n = __import__("ctypes").c_int(0)
    assert ref._L.ref_extract(ref._h, None, 0, 0, 0, 0, 1000, None, None, 0, n) == -1


def test_quadtree_alone(oracle_mod):
    """DistributeOctTree on synthetic key sets (dense, clustered, tiny, more roots than keys): reference == restatement."""
    rng = np.random.default_rng(5)
    for n, w, h, N in [(3000, 720, 448, 217), (800, 720, 448, 217), (150, 178, 102, 60), (5000, 1209, 344, 434),
                       (40, 720, 448, 217), (2000, 300, 300, 1086), (1, 720, 448, 10), (3, 100, 133, 5), (60000, 720, 448, 1086)]:
        x, y = rng.integers(0, w, n), rng.integers(0, h, n)
        xy = np.unique(np.stack([x, y], 1), axis=0)
        xy = xy[np.lexsort((xy[:, 0], xy[:, 1]))]
        # few distinct responses => many compareNodes ties and many equal-response picks inside a node
        keys = np.concatenate([xy, rng.integers(7, 12, (len(xy), 1))], 1).astype(np.int32)
        a = oracle_mod.distribute(keys, 16, 16 + w, 16, 16 + h, N)
        b = oref.distribute(keys, 16, 16 + w, 16, 16 + h, N)
        assert np.array_equal(a, b), (n, w, h, N)


def test_reference_release_flags_fma(oracle_mod):
    """The reference's Release flags (-O3 -march=native, CMakeLists.txt:10-13) let GCC contract x*b + y*a into FMAs.
    That may move a BRIEF sample at an exact cvRound half-way point and nothing else: same keypoints and angles, and
    at least 99.9 % identical descriptors (BASELINE.json's tolerance)."""
    if not oref.available(fma=True):
        pytest.skip("FMA build absent")
    tot = same = 0
    for seed, (w, h, nf) in enumerate([(752, 480, 1000), (1241, 376, 2000), (752, 480, 1000)]):
        img = synth.frame(w, h, 20 + seed)
        a = oracle_mod.Extractor(nf, 1.2, 8, 20, 7)(img)
        b = oref.Extractor(nf, 1.2, 8, 20, 7, fma=True)(img)
        assert a[0] == b[0]
        assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))
        tot += len(a[2])
        same += int((a[2] == b[2]).all(axis=1).sum())
    assert same / tot >= 0.999, (same, tot)
    print(f"descriptors identical under the reference's FMA build: {same}/{tot}")
