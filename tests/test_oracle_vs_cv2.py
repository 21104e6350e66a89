"""CPU: the oracle against LIVE cv2 (when importable) on random inputs — wider than the committed fixtures."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
from mam3slam_b200 import synth   # noqa: E402


@pytest.mark.parametrize("seed", range(6))
def test_resize_random_shapes(oracle_mod, seed):
    rng = np.random.default_rng(seed)
    sw, sh = int(rng.integers(40, 400)), int(rng.integers(40, 300))
    dw, dh = int(round(sw / 1.2)), int(round(sh / 1.2))
    src = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
    assert np.array_equal(oracle_mod.resize_linear(src, dw, dh), cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR))


def test_resize_euroc_chain(oracle_mod):
    img = synth.frame(752, 480, 0)
    sizes = [(627, 400), (522, 333), (435, 278), (363, 231), (302, 193), (252, 161), (210, 134)]
    a = b = img
    for w, h in sizes:
        a = oracle_mod.resize_linear(a, w, h)
        b = cv2.resize(b, (w, h), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(a, b), (w, h)


def test_exact_half_uses_area_path(oracle_mod):
    rng = np.random.default_rng(1)
    src = rng.integers(0, 256, (96, 128), dtype=np.uint8)
    assert np.array_equal(oracle_mod.resize_linear(src, 64, 48), cv2.resize(src, (64, 48), interpolation=cv2.INTER_LINEAR))


@pytest.mark.parametrize("shape", [(64, 80), (7, 7), (9, 200), (133, 4)])
def test_gaussian_shapes(oracle_mod, shape):
    rng = np.random.default_rng(shape[0])
    src = rng.integers(0, 256, shape, dtype=np.uint8)
    want = cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    assert np.array_equal(oracle_mod.gaussian7(src), want)


@pytest.mark.parametrize("th", [20, 7, 1, 60])
def test_fast_cells_random(oracle_mod, th):
    rng = np.random.default_rng(th)
    det = cv2.FastFeatureDetector_create(th, True)
    tex = synth.frame(300, 200, 2)
    n = 0
    for i in range(30):
        w, h = int(rng.integers(7, 48)), int(rng.integers(7, 48))
        if i % 2:
            cell = rng.integers(0, 256, (h, w), dtype=np.uint8)
        else:
            x, y = int(rng.integers(0, 300 - w)), int(rng.integers(0, 200 - h))
            cell = np.ascontiguousarray(tex[y:y + h, x:x + w])
        want = np.array([(k.pt[0], k.pt[1], k.response) for k in det.detect(cell)], np.int32).reshape(-1, 3)
        assert np.array_equal(oracle_mod.fast9_nms(cell, th), want)
        n += len(want)
    assert n > 0


def test_full_frame_stages_vs_cv2(oracle_mod):
    from oracle.py import cv2_pipeline as cvp
    img = synth.frame(752, 480, 1)
    ex = oracle_mod.Extractor(1000, 1.2, 8, 20, 7)
    ex(img)
    for l, (padded, roi) in enumerate(cvp.pyramid(img, 1.2, 8)):
        assert np.array_equal(ex.level_pixels(l, "padded"), padded)
        cand, _, _ = cvp.fast_cells(np.ascontiguousarray(roi), 20, 7)
        assert np.array_equal(ex.level_points(l)[:, :3], cand)
        assert np.array_equal(ex.level_pixels(l, "blurred"), cvp.blur(roi))
