"""GPU parity of the Frame-side entry points through the C ABI: undistortion bit-exact against the cv2 golden
vectors and the oracle, the DBoW2 transform (word / node ids, BowVector doubles, FeatureVector CSR) identical to the oracle's
restatement, and the vocabulary validation."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from mam3slam_b200 import ORBmatcher, KEYPOINT_DTYPE
from mam3slam_b200._lib import NativeError
from oracle import matcher as om
from tests import matcher_scenarios as sc

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "cv2_undistort.npz")


@pytest.fixture(scope="module")
def M():
    m = ORBmatcher(0.9, True)
    yield m
    m.close()


def test_undistort_points_bit_exact_vs_cv2_golden(M):
    g = np.load(GOLD)
    for cam in sorted({k.rsplit("_", 1)[0] for k in g.files}):
        got = M.UndistortPoints(g[cam + "_xy"], g[cam + "_K"], g[cam + "_dist"])
        assert np.array_equal(got.view(np.uint32), g[cam + "_out"].view(np.uint32)), cam


def test_undistort_keypoints_device_batch(M):
    rng = np.random.default_rng(2)
    nframes, stride = 5, 700
    counts = np.array([700, 0, 1, 333, 699], np.int32)
    kps = np.zeros((nframes, stride), KEYPOINT_DTYPE)
    kps["x"] = rng.uniform(0, 752, kps.shape)
    kps["y"] = rng.uniform(0, 480, kps.shape)
    for f in ("size", "angle", "response"):
        kps[f] = rng.uniform(0, 360, kps.shape)
    kps["octave"] = rng.integers(0, 8, kps.shape)
    kps["class_id"] = -1
    K4, dist = (458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05)
    d_in = torch.from_numpy(kps.view(np.uint8).reshape(nframes, -1)).cuda()
    d_out = torch.full_like(d_in, 0xEE)
    d_cnt = torch.from_numpy(counts).cuda()
    M.undistort_keypoints_device(nframes, d_in.data_ptr(), d_cnt.data_ptr(), stride, K4, dist, d_out.data_ptr())
    torch.cuda.synchronize()
    out = d_out.cpu().numpy().view(KEYPOINT_DTYPE).reshape(nframes, stride)
    for f in range(nframes):
        n = counts[f]
        want = om.undistort_points(np.stack([kps["x"][f, :n], kps["y"][f, :n]], 1), K4, dist)
        assert np.array_equal(out["x"][f, :n], want[:, 0]) and np.array_equal(out["y"][f, :n], want[:, 1])
        for fld in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(out[fld][f, :n], kps[fld][f, :n])
        assert (out.view(np.uint8).reshape(nframes, stride, 28)[f, n:] == 0xEE).all()      # nothing past the count


def test_undistort_rejects_bad_arguments(M):
    xy = np.zeros((4, 2), np.float32)
    with pytest.raises(NativeError):
        M.UndistortPoints(xy, (1, 1, 0, 0), (0.1, 0.0, 0.0))                  # 3 coefficients
    with pytest.raises(NativeError):
        M.UndistortPoints(xy, (0, 1, 0, 0), (0.1, 0.0, 0.0, 0.0))             # fx = 0
    assert M.UndistortPoints(np.zeros((0, 2), np.float32), (1, 1, 0, 0), (0.1, 0, 0, 0)).shape == (0, 2)


def _voc(M, v):
    return M.vocabulary(v["L"], v["child_off"], v["children"], v["desc"], v["word_id"], v["weight"], v["weighting"], v["norm"])


@pytest.mark.parametrize("weighting,norm", [(0, 1), (1, 0), (2, 2), (3, 1), (0, 2), (1, 1)])
def test_bow_transform_matches_oracle(M, weighting, norm):
    v = sc.make_vocab(k=10, L=4, seed=3, weighting=weighting, norm=norm)
    V = _voc(M, v)
    desc = sc.vocab_features(v, 1500)
    for levelsup in (0, 2, 4, 6):
        (bw, bv), (fn, fo, fi), w, nd = V.transform(desc, levelsup)
        r = om.bow_transform(v, desc, levelsup)
        assert np.array_equal(w, r["word"]) and np.array_equal(nd, r["node"])
        assert np.array_equal(bw, r["bow_word"])
        assert np.array_equal(bv.view(np.uint64), r["bow_value"].view(np.uint64))       # doubles bit-identical
        assert np.array_equal(fn, r["fv_node"]) and np.array_equal(fo, r["fv_off"]) and np.array_equal(fi, r["fv_idx"])
    assert (w < 0).any()
    V.close()


def test_bow_transform_orbvoc_shape_and_real_descriptors(M):
    """k = 10, L = 5 (111 k nodes; the real ORBvoc is k = 10, L = 6) with the descriptors of an extracted frame, then the
    FeatureVector goes straight into SearchByBoW."""
    v = sc.make_vocab(k=10, L=5, seed=9, ragged=False, stop_frac=0.01)
    V = _voc(M, v)
    (k1, d1), (k2, d2), bounds, _ = sc.frame_pair(752, 480, 1000, 0)
    outs = []
    for d in (d1, d2):
        (bw, bv), fv, w, nd = V.transform(d, 3)
        r = om.bow_transform(v, d, 3)
        assert np.array_equal(w, r["word"]) and np.array_equal(nd, r["node"])
        assert np.array_equal(bv.view(np.uint64), r["bow_value"].view(np.uint64))
        assert np.array_equal(fv[2], r["fv_idx"])
        outs.append(fv)
    F1, F2 = M.frame(k1, d1, bounds), M.frame(k2, d2, bounds)
    elig = np.ones(len(k1), np.uint8)
    got = M.SearchByBoW_KF_Frame(F1, outs[0], elig, F2, outs[1])
    want = om.search_bow_kf_frame(k1, d1, outs[0], elig, k2, d2, outs[1], 0.9, True)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])
    V.close()


def test_bow_transform_device_batch(M):
    v = sc.make_vocab(k=8, L=4, seed=21)
    V = _voc(M, v)
    nframes, stride = 4, 900
    counts = np.array([900, 0, 417, 1], np.int32)
    desc = np.stack([sc.vocab_features(v, stride, seed=30 + f) for f in range(nframes)])
    d_desc = torch.from_numpy(desc).cuda()
    d_cnt = torch.from_numpy(counts).cuda()
    d_w = torch.full((nframes, stride), -77, dtype=torch.int32, device="cuda")
    d_n = torch.full((nframes, stride), -77, dtype=torch.int32, device="cuda")
    V.transform_device(nframes, d_desc.data_ptr(), d_cnt.data_ptr(), stride, 2, d_w.data_ptr(), d_n.data_ptr())
    torch.cuda.synchronize()
    w, nd = d_w.cpu().numpy(), d_n.cpu().numpy()
    for f in range(nframes):
        n = counts[f]
        r = om.bow_transform(v, desc[f, :n], 2)
        assert np.array_equal(w[f, :n], r["word"]) and np.array_equal(nd[f, :n], r["node"])
        assert (w[f, n:] == -77).all() and (nd[f, n:] == -77).all()
    V.close()


def test_bow_empty_and_invalid_vocabulary(M):
    v = sc.make_vocab(k=4, L=2, seed=1)
    V = _voc(M, v)
    (bw, bv), (fn, fo, fi), w, nd = V.transform(np.zeros((0, 32), np.uint8), 1)
    assert len(bw) == 0 and len(fn) == 0 and fo.tolist() == [0]
    V.close()
    bad = dict(v)
    ch = v["children"].copy()
    ch[3] = ch[2]                                       # a node reached twice
    bad["children"] = ch
    with pytest.raises(NativeError):
        _voc(M, bad)
    bad = dict(v)
    ch = v["children"].copy()
    ch[0] = len(v["child_off"]) + 5                     # out of range
    bad["children"] = ch
    with pytest.raises(NativeError):
        _voc(M, bad)
    bad = dict(v)
    co = v["child_off"].copy()
    co[2] = co[1] - 1                                   # not monotone
    bad["child_off"] = co
    with pytest.raises(NativeError):
        _voc(M, bad)
