"""Frame::ComputeStereoMatches (reference src/Frame.cc:811-981).
CPU: the oracle restatement against an independent numpy restatement (row-band candidate set, first-minimum Hamming,
11-shift SAD, parabola, median filter).  GPU: orbx_stereo_matches on two device-resident extractor results, bit-exact."""
import numpy as np
import pytest

import oracle
from oracle import matcher as om
from mam3slam_b200 import synth
from tests.matcher_scenarios import to_kp

MB, MBF = 0.11, 47.9          # EuRoC-like baseline (m) and baseline * fx
f32 = np.float32


def _oracle_side(w, h, nfeat, seed):
    left, right = synth.stereo_pair(w, h, seed)
    out = []
    for im in (left, right):
        ex = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
        _, k, d = ex(im)
        pyr = [ex.level_pixels(l) for l in range(8)]
        out.append((to_kp(k), d, pyr, ex.tables()))
    return left, right, out


def _numpy_stereo(kL, dL, kR, dR, pyrL, pyrR, scale, inv_scale):
    N = len(kL)
    ur = np.full(N, -1, f32); dp = np.full(N, -1, f32)
    maxD = f32(MBF) / f32(MB)
    r = f32(2) * scale[kR["octave"]]
    maxr = np.ceil(kR["y"] + r).astype(int); minr = np.floor(kR["y"] - r).astype(int)
    pairs = []
    for i in range(N):
        lv, uL, row = int(kL["octave"][i]), kL["x"][i], int(kL["y"][i])
        m = (minr <= row) & (row <= maxr) & (np.abs(kR["octave"] - lv) <= 1) & (kR["x"] >= uL - maxD) & (kR["x"] <= uL)
        c = np.nonzero(m)[0]
        if len(c) == 0:
            continue
        dist = np.unpackbits(dR[c] ^ dL[i][None, :], axis=1).sum(1)
        j = int(np.argmin(dist))
        if dist[j] >= 75:
            continue
        sf = inv_scale[lv]
        rnd = lambda x: f32(np.floor(abs(x) + f32(0.5)) * np.sign(x))         # round(): half away from zero
        su, sv, sr = int(rnd(kL["x"][i] * sf)), int(rnd(kL["y"][i] * sf)), int(rnd(kR["x"][c[j]] * sf))
        if sr < 0 or sr + 11 >= pyrR[lv].shape[1]:
            continue
        IL = pyrL[lv][sv - 5:sv + 6, su - 5:su + 6].astype(np.int64)
        sads = np.array([np.abs(IL - pyrR[lv][sv - 5:sv + 6, sr + inc - 5:sr + inc + 6]).sum() for inc in range(-5, 6)])
        b = int(np.argmin(sads))
        if b in (0, 10):
            continue
        d1, d2, d3 = f32(sads[b - 1]), f32(sads[b]), f32(sads[b + 1])
        with np.errstate(divide="ignore", invalid="ignore"):
            delta = (d1 - d3) / (f32(2) * (d1 + d3 - f32(2) * d2))
        if delta < -1 or delta > 1 or np.isnan(delta):
            if not np.isnan(delta):
                continue
        best_u = scale[lv] * (f32(sr) + f32(b - 5) + delta)
        disp = uL - best_u
        if disp >= 0 and disp < maxD:
            if disp <= 0:
                disp = f32(0.01); best_u = f32(np.float64(uL) - 0.01)
            dp[i] = f32(MBF) / disp; ur[i] = best_u
            pairs.append((int(sads[b]), i))
    if pairs:
        pairs.sort()
        th = f32(f32(1.5) * f32(1.4)) * f32(pairs[len(pairs) // 2][0])
        for s, i in pairs:
            if not f32(s) < th:
                ur[i] = -1; dp[i] = -1
    return ur, dp


def test_oracle_stereo_vs_numpy():
    _, _, ((kL, dL, pL, tabs), (kR, dR, pR, _)) = _oracle_side(376, 240, 500, 3)
    kept, ur, dp = om.stereo_matches(kL, dL, kR, dR, pL, pR, tabs["scale"], tabs["inv_scale"], MB, MBF)
    wur, wdp = _numpy_stereo(kL, dL, kR, dR, pL, pR, tabs["scale"].astype(f32), tabs["inv_scale"].astype(f32))
    assert np.array_equal(ur, wur) and np.array_equal(dp, wdp)
    assert kept == (ur >= 0).sum() and kept > 50
    # the planted disparity (17.3 px) is recovered to sub-pixel accuracy by most of the kept matches
    disp = kL["x"][ur >= 0] - ur[ur >= 0]
    assert np.median(np.abs(disp - 17.3)) < 0.3


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(752, 480, 1000, 0), (1241, 376, 2000, 1)], ids=["euroc", "kitti"])
def test_stereo_matches_gpu(shape):
    from mam3slam_b200 import ORBextractor
    from mam3slam_b200.extractor import ComputeStereoMatches
    w, h, nfeat, seed = shape
    left, right, ((kL, dL, pL, tabs), (kR, dR, pR, _)) = _oracle_side(w, h, nfeat, seed)
    okept, our, odp = om.stereo_matches(kL, dL, kR, dR, pL, pR, tabs["scale"], tabs["inv_scale"], MB, MBF)
    exL, exR = ORBextractor(nfeat, 1.2, 8, 20, 7), ORBextractor(nfeat, 1.2, 8, 20, 7)
    _, gkL, gdL = exL(left)
    _, gkR, gdR = exR(right)
    assert np.array_equal(gkL["x"], kL["x"]) and np.array_equal(gdR, dR)
    kept, ur, dp = ComputeStereoMatches(exL, exR, MB, MBF)
    assert kept == okept and np.array_equal(ur, our) and np.array_equal(dp, odp)
    assert kept > 200
    # an image pair without texture has no stereo matches at all
    flat = np.full((h, w), 90, np.uint8)
    exL(flat); exR(flat)
    kept, ur, dp = ComputeStereoMatches(exL, exR, MB, MBF)
    assert kept == 0 and len(ur) == 0
    exL.close(); exR.close()
