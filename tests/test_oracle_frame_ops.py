"""CPU: the oracle's restatements of the Frame-side steps next to the matcher —
* cv::undistortPoints as Frame::UndistortKeyPoints calls it, pinned bit-exact against cv2 golden vectors
  (tests/golden/cv2_undistort.npz, made by tools/gen_golden_undistort.py);
* DBoW2 transform (TemplatedVocabulary.h:1127-1260), against a second, dict-based restatement written here.  DBoW2's own
  sources need OpenCV headers and the vocabulary file is absent from the checkout, so this path has no reference-run
  fixture: the transform's parity is "unpinned" beyond the two independent restatements agreeing."""
import os

import numpy as np
import pytest

from oracle import matcher as om
from tests import matcher_scenarios as sc

GOLD = os.path.join(os.path.dirname(__file__), "golden", "cv2_undistort.npz")


def test_undistort_points_bit_exact_vs_cv2():
    g = np.load(GOLD)
    cams = sorted({k.rsplit("_", 1)[0] for k in g.files})
    assert len(cams) == 6
    for cam in cams:
        got = om.undistort_points(g[cam + "_xy"], g[cam + "_K"], g[cam + "_dist"])
        assert np.array_equal(got.view(np.uint32), g[cam + "_out"].view(np.uint32)), cam


def test_undistort_zero_distortion_is_identity_up_to_rounding():
    xy = np.array([[10, 20], [300.5, 17.25], [751, 479]], np.float32)
    out = om.undistort_points(xy, (458.654, 457.296, 367.215, 248.375), (0, 0, 0, 0))
    assert np.abs(out - xy).max() < 1e-4


def py_transform(voc, desc, levelsup):
    """second restatement: literal dict / list walk of TemplatedVocabulary::transform"""
    co, ch, nd, wi, we = voc["child_off"], voc["children"], voc["desc"], voc["word_id"], voc["weight"]
    bow, fv, words, nodes = {}, {}, [], []
    nid_level = voc["L"] - levelsup
    for i, f in enumerate(desc):
        nid = 0
        final, level = 0, 0
        while True:
            level += 1
            kids = ch[co[final]:co[final + 1]]
            final = int(kids[0])
            best = int(np.unpackbits(f ^ nd[final]).sum())
            for c in kids[1:]:
                d = int(np.unpackbits(f ^ nd[c]).sum())
                if d < best:
                    best, final = d, int(c)
            if level == nid_level:
                nid = final
            if co[final + 1] == co[final]:
                break
        w = float(we[final])
        word = int(wi[final])
        nodes.append(nid)
        if w > 0:
            words.append(word)
            if voc["weighting"] in (0, 1):
                bow[word] = bow.get(word, 0.0) + w if word in bow else w
            else:
                bow.setdefault(word, w)
            fv.setdefault(nid, []).append(i)
        else:
            words.append(-1 - word)
    keys = sorted(bow)
    vals = [bow[k] for k in keys]
    if voc["weighting"] in (0, 1) and keys and voc["norm"] == 0:
        vals = [v / float(len(keys)) for v in vals]
    if voc["norm"]:
        norm = 0.0
        for v in vals:
            norm += abs(v) if voc["norm"] == 1 else v * v
        if voc["norm"] == 2:
            norm = float(np.sqrt(norm))
        if norm > 0:
            vals = [v / norm for v in vals]
    return keys, vals, {k: fv[k] for k in sorted(fv)}, words, nodes


@pytest.mark.parametrize("weighting,norm", [(0, 1), (1, 0), (2, 2), (3, 1)])
def test_bow_transform_vs_second_restatement(weighting, norm):
    voc = sc.make_vocab(k=6, L=3, seed=5, stop_frac=0.15, weighting=weighting, norm=norm)
    desc = sc.vocab_features(voc, 300)
    for levelsup in (0, 1, 2, 3, 5):
        r = om.bow_transform(voc, desc, levelsup)
        keys, vals, fv, words, nodes = py_transform(voc, desc, levelsup)
        assert r["word"].tolist() == words and r["node"].tolist() == nodes
        assert r["bow_word"].tolist() == keys
        assert r["bow_value"].tolist() == vals                      # doubles, same operation order -> identical
        assert r["fv_node"].tolist() == list(fv)
        for j, k in enumerate(fv):
            assert r["fv_idx"][r["fv_off"][j]:r["fv_off"][j + 1]].tolist() == fv[k]
    assert (np.array(words) < 0).any()                              # stopped words occur in the scenario


def test_bow_transform_properties():
    voc = sc.make_vocab(k=10, L=4, seed=3)
    desc = sc.vocab_features(voc, 1000)
    r = om.bow_transform(voc, desc, 4)
    assert (r["node"] == 0).all() and list(r["fv_node"]) == [0]     # levelsup >= L: everything under the root
    r = om.bow_transform(voc, desc, 2)
    kept = np.nonzero(r["word"] >= 0)[0]
    assert sorted(r["fv_idx"].tolist()) == kept.tolist()            # every unstopped feature exactly once
    assert abs(r["bow_value"].sum() - 1.0) < 1e-12                  # L1-normalised
    assert (np.diff(r["bow_word"]) > 0).all() and (np.diff(r["fv_node"]) > 0).all()
    assert r["word"][5] == r["word"][23] == r["word"][31]
    # a leaf descriptor itself descends to its own word unless an earlier duplicate sibling wins the tie
    leaves = np.nonzero(voc["word_id"] >= 0)[0][:200]
    rr = om.bow_transform(voc, voc["desc"][leaves], 0)
    same = np.abs(np.where(rr["word"] < 0, -1 - rr["word"], rr["word"])) == voc["word_id"][leaves]
    assert same.mean() > 0.8
