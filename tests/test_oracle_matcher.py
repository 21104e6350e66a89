"""CPU: the matcher oracle (oracle/cpp/orc_matcher.cc) against independent numpy restatements where one exists
(DescriptorDistance, k-NN, the frame grid, GetFeaturesInArea, ComputeThreeMaxima), the sequential-semantics properties of
the searches, and the host-only entry points of the C ABI (orbm_distance, orbm_knn_merge)."""
import numpy as np
import pytest

from oracle import matcher as om
from tests import matcher_scenarios as sc


def _popdist(a, b):
    return np.unpackbits(a ^ b, axis=-1).sum(-1)


def test_descriptor_distance_vs_numpy():
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    b[:10] = a[:10]
    b[10:20] = ~a[10:20]
    want = _popdist(a, b)
    assert [om.descriptor_distance(a[i], b[i]) for i in range(300)] == want.tolist()
    assert want[:10].tolist() == [0] * 10 and want[10:20].tolist() == [256] * 10


def test_abi_distance_and_merge_host_side():
    from mam3slam_b200.matcher import ORBmatcher
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    assert [ORBmatcher.DescriptorDistance(a[i], b[i]) for i in range(50)] == _popdist(a, b).tolist()
    # merge of sharded top-k lists == top-k of the whole (ties -> lowest global row)
    q, db = sc.knn_data(4000, 40)
    k = 4
    full_d, full_i = om.knn(q, db, k)
    parts_d, parts_i = [], []
    for s in range(4):
        d, i = om.knn(q, db[s * 1000:(s + 1) * 1000], k)
        parts_d.append(d); parts_i.append(i + s * 1000)
    md, mi = ORBmatcher.knn_merge(np.stack(parts_d), np.stack(parts_i))
    assert np.array_equal(md, full_d) and np.array_equal(mi, full_i)


def test_knn_vs_numpy_with_ties():
    q, db = sc.knn_data(3000, 64)
    d, i = om.knn(q, db, 5, threads=4)
    D = _popdist(q[:, None, :], db[None, :, :])
    order = np.lexsort((np.broadcast_to(np.arange(len(db)), D.shape), D), axis=1)[:, :5]
    assert np.array_equal(i, order)
    assert np.array_equal(d, np.take_along_axis(D, order, 1))
    assert (d[:, 0] <= 40).all()          # the planted neighbours are found


def test_grid_and_features_in_area_vs_numpy():
    (k1, d1), _, bounds, _ = sc.frame_pair()
    off, idx = om.grid_build(k1, bounds)
    invw = np.float32(64) / np.float32(bounds[2] - bounds[0])
    invh = np.float32(48) / np.float32(bounds[3] - bounds[1])
    px = np.floor((k1["x"] - np.float32(bounds[0])) * invw + np.float32(0.5)).astype(int)     # round() for x >= 0
    py = np.floor((k1["y"] - np.float32(bounds[1])) * invh + np.float32(0.5)).astype(int)
    for c in range(64 * 48):
        want = np.nonzero((px == c // 48) & (py == c % 48))[0]
        assert np.array_equal(idx[off[c]:off[c + 1]], want)
    rng = np.random.default_rng(2)
    for _ in range(50):
        x, y, r = float(rng.uniform(-20, 780)), float(rng.uniform(-20, 500)), float(rng.uniform(1, 120))
        lo, hi = int(rng.integers(-1, 7)), int(rng.integers(-1, 8))
        got = om.features_in_area(k1, bounds, x, y, r, lo, hi)
        m = (np.abs(k1["x"] - np.float32(x)) < np.float32(r)) & (np.abs(k1["y"] - np.float32(y)) < np.float32(r))
        if lo > 0 or hi >= 0:
            m &= k1["octave"] >= lo
            if hi >= 0:
                m &= k1["octave"] <= hi
        m &= (px >= 0) & (px < 64) & (py >= 0) & (py < 48)
        assert sorted(got.tolist()) == np.nonzero(m)[0].tolist()
        # order: cell-x outer, cell-y inner, insertion order inside a cell
        cells = (px[got] * 48 + py[got])
        assert np.all(np.diff(cells) >= 0)
        for c in np.unique(cells):
            assert np.all(np.diff(got[cells == c]) > 0)


def test_three_maxima_cases():
    assert om.three_maxima([0] * 30) == (-1, -1, -1)
    h = [0] * 30; h[3] = 100; h[7] = 9; h[9] = 50
    assert om.three_maxima(h) == (3, 9, -1)          # 9 < 0.1*100 -> third dropped
    h[7] = 10
    assert om.three_maxima(h) == (3, 9, 7)
    h = [0] * 30; h[1] = 100; h[2] = 5
    assert om.three_maxima(h) == (1, -1, -1)
    h = [4] * 30
    assert om.three_maxima(h) == (0, 1, 2)           # first seen wins ties
    assert om.rot_bin(10.0, 350.0) == 1 and om.rot_bin(0.0, 0.0) == 0 and om.rot_bin(359.0, 0.0) == 12


def test_search_frame_sequential_semantics():
    """M2 properties: a keypoint is claimed at most once, claimed keypoints were free, every match is within TH_HIGH,
    and re-running with the result as `occupied` finds nothing new for the same queries' winners."""
    (k1, d1), (k2, d2), bounds, tabs = sc.frame_pair()
    n1 = len(k1)
    occ = np.zeros(len(k2), np.uint8)
    occ[::7] = 1
    nm, km = om.search_frame(k2, d2, bounds, occ, tabs["scale"], np.ones(n1, np.uint8), k1["x"], k1["y"], k1["octave"],
                             k1["angle"], d1, 15.0, 100, True)
    assert nm == (km >= 0).sum() and nm > 200
    m = np.nonzero(km >= 0)[0]
    assert not occ[m].any()
    assert len(np.unique(km[m])) == len(m)
    dist = _popdist(d2[m], d1[km[m]])
    assert (dist <= 100).all()
    assert (np.abs(k2["x"][m] - k1["x"][km[m]]) < 15 * tabs["scale"][k1["octave"][km[m]]]).all()
    # without the rotation check there are at least as many matches
    nm2, _ = om.search_frame(k2, d2, bounds, occ, tabs["scale"], np.ones(n1, np.uint8), k1["x"], k1["y"], k1["octave"],
                             k1["angle"], d1, 15.0, 100, False)
    assert nm2 >= nm


def test_distinctive_descriptor_vs_numpy():
    rng = np.random.default_rng(3)
    for n in (1, 2, 3, 8, 21):
        d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        D = _popdist(d[:, None, :], d[None, :, :])
        med = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]
        assert om.distinctive_descriptor(d) == int(np.argmin(med))


def test_search_by_sim3_vs_numpy():
    """M8: each direction = first minimum over the window candidates of levels [l-1, l], accepted at <= TH_HIGH; result =
    mutual agreement.  Restated with numpy on top of the (numpy-pinned) GetFeaturesInArea."""
    (k1, d1), (k2, d2), bounds, tabs = sc.frame_pair()
    rng = np.random.default_rng(29)
    n1, n2 = len(k1), len(k2)
    l12 = np.clip(k1["octave"] + rng.integers(0, 2, n1), 0, 7).astype(np.int32)
    l21 = np.clip(k2["octave"] + rng.integers(0, 2, n2), 0, 7).astype(np.int32)
    v1 = (rng.random(n1) < 0.9).astype(np.uint8)
    v2 = (rng.random(n2) < 0.9).astype(np.uint8)
    th = 9.0

    def one_way(kq, dq, valid, lev, kt, dt):
        out = np.full(len(kq), -1)
        for i in range(len(kq)):
            if not valid[i]:
                continue
            r = np.float32(th) * tabs["scale"][lev[i]]
            cand = om.features_in_area(kt, bounds, float(kq["x"][i]), float(kq["y"][i]), float(r), -1, -1)
            cand = cand[(kt["octave"][cand] >= lev[i] - 1) & (kt["octave"][cand] <= lev[i])]
            if len(cand) == 0:
                continue
            dist = _popdist(dt[cand], dq[i][None, :])
            j = int(np.argmin(dist))                  # first minimum
            if dist[j] <= 100:
                out[i] = cand[j]
        return out

    m1 = one_way(k1, d1, v1, l12, k2, d2)
    m2 = one_way(k2, d2, v2, l21, k1, d1)
    want = np.array([m1[i] if m1[i] >= 0 and m2[m1[i]] == i else -1 for i in range(n1)])
    nf, m12 = om.search_by_sim3(k1, d1, bounds, k2, d2, bounds, tabs["scale"], tabs["scale"], v1, k1["x"], k1["y"], l12, d1, v2,
                                k2["x"], k2["y"], l21, d2, th)
    assert np.array_equal(m12, want) and nf == (want >= 0).sum() and nf > 20
