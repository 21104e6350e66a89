"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes bindings of oracle/cpp/orc_matcher.cc (CPU restatement of the reference's
ORBmatcher searches over plain arrays).  Keypoints are numpy structured arrays with cv::KeyPoint's 28-byte layout."""
import ctypes as C
import numpy as np

from . import lib, _p

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30


def _c(a, dt):
    return np.ascontiguousarray(a, dt)


def _kp(k):
    k = np.ascontiguousarray(k)
    assert k.dtype.itemsize == 28
    return k


def _bounds(b):
    return _c(b, np.float32)


def _L():
    L = lib()
    if not getattr(L, "_matcher_ready", False):
        L.orc_knn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_long, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_features_in_area.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int,
                                           C.c_int, C.c_void_p]
        L.orc_rot_bin.argtypes = [C.c_float, C.c_float]
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        L.orc_search_local_map.argtypes = [i, vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, f, f, vp, vp]
        L.orc_search_frame.argtypes = [i, vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, f, i, i, vp, vp]
        L.orc_search_sim3.argtypes = [i, vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, i, f, vp]
        L.orc_search_initialization.argtypes = [i, vp, vp, i, vp, vp, vp, vp, i, f, i, vp]
        L.orc_search_bow_kf_frame.argtypes = [i, vp, vp, i, vp, vp, vp, vp, i, vp, vp, i, vp, vp, vp, f, i, vp]
        L.orc_search_bow_kf_kf.argtypes = [i, vp, vp, i, vp, vp, vp, vp, i, vp, vp, i, vp, vp, vp, vp, f, i, vp]
        L.orc_search_triangulation.argtypes = [i, vp, vp, i, vp, vp, vp, vp, i, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, vp,
                                               i, i, vp]
        L.orc_fuse_candidates.argtypes = [i, vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, f, i, vp, vp]
        L.orc_search_by_sim3.argtypes = [i, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, f, vp]
        L.orc_stereo_matches.argtypes = [i, vp, vp, i, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, f, f, vp, vp]
        L.orc_distinctive_descriptor.argtypes = [vp, i]
        L.orc_undistort_points.argtypes = [vp, i, vp, vp, i, vp]
        L.orc_bow_transform.argtypes = [i, i, vp, vp, vp, vp, vp, i, i, vp, i, i, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L._matcher_ready = True
    return L


def descriptor_distance(a, b):
    a, b = _c(a, np.uint8), _c(b, np.uint8)
    return int(_L().orc_descriptor_distance(_p(a), _p(b)))


def grid_build(kps, bounds):
    kps, b = _kp(kps), _bounds(bounds)
    off = np.empty(64 * 48 + 1, np.int32)
    idx = np.empty(max(len(kps), 1), np.int32)
    _L().orc_grid_build(len(kps), _p(kps), _p(b), _p(off), _p(idx))
    return off, idx[:off[-1]].copy()


def features_in_area(kps, bounds, x, y, r, min_level=-1, max_level=-1):
    kps, b = _kp(kps), _bounds(bounds)
    out = np.empty(max(len(kps), 1), np.int32)
    n = _L().orc_features_in_area(len(kps), _p(kps), _p(b), float(x), float(y), float(r), int(min_level), int(max_level), _p(out))
    return out[:n].copy()


def three_maxima(sizes):
    s = _c(sizes, np.int32)
    ind = np.empty(3, np.int32)
    _L().orc_three_maxima(_p(s), len(s), _p(ind))
    return tuple(int(v) for v in ind)


def rot_bin(a1, a2):
    return int(_L().orc_rot_bin(float(a1), float(a2)))


def search_local_map(kps, desc, bounds, occupied, scale_factors, valid, proj_x, proj_y, level, view_cos, qdesc, th, nnratio,
                     q_occupies=None):
    """M1.  Returns (nmatches, kpMatch[n], occupied_after[n]).  q_occupies[q] = the query's map point has Observations() > 0
    (src/ORBmatcher.cc:92-94): only then does its match block the keypoint for later queries; None = always."""
    qo = None if q_occupies is None else _c(q_occupies, np.uint8)
    kps, desc, b = _kp(kps), _c(desc, np.uint8), _bounds(bounds)
    occ = _c(occupied, np.uint8).copy()
    sf = _c(scale_factors, np.float32)
    valid, px, py = _c(valid, np.uint8), _c(proj_x, np.float32), _c(proj_y, np.float32)
    level, vc, qd = _c(level, np.int32), _c(view_cos, np.float32), _c(qdesc, np.uint8)
    km = np.full(len(kps), -1, np.int32)
    n = _L().orc_search_local_map(len(kps), _p(kps), _p(desc), _p(b), _p(occ), _p(sf), len(valid), _p(valid), _p(px), _p(py),
                                  _p(level), _p(vc), _p(qd), float(th), float(nnratio), _p(km), None if qo is None else _p(qo))
    return n, km, occ


def search_frame(kps, desc, bounds, occupied, scale_factors, valid, u, v, octave, qangle, qdesc, th, th_dist=TH_HIGH,
                 check_ori=True, q_occupies=None):
    """M2 / M3.  Returns (nmatches, kpMatch[n]).  q_occupies: as in search_local_map (src/ORBmatcher.cc:1745-1747)."""
    qo = None if q_occupies is None else _c(q_occupies, np.uint8)
    kps, desc, b = _kp(kps), _c(desc, np.uint8), _bounds(bounds)
    occ = _c(occupied, np.uint8).copy()
    sf = _c(scale_factors, np.float32)
    valid, u, v = _c(valid, np.uint8), _c(u, np.float32), _c(v, np.float32)
    octave, qa, qd = _c(octave, np.int32), _c(qangle, np.float32), _c(qdesc, np.uint8)
    km = np.full(len(kps), -1, np.int32)
    n = _L().orc_search_frame(len(kps), _p(kps), _p(desc), _p(b), _p(occ), _p(sf), len(valid), _p(valid), _p(u), _p(v),
                              _p(octave), _p(qa), _p(qd), float(th), int(th_dist), int(check_ori), _p(km),
                              None if qo is None else _p(qo))
    return n, km


def search_sim3(kps, desc, bounds, matched, scale_factors, valid, u, v, level, qdesc, th, ratio_hamming):
    """M7.  Returns (nmatches, kpMatch[n])."""
    kps, desc, b = _kp(kps), _c(desc, np.uint8), _bounds(bounds)
    m = _c(matched, np.uint8).copy()
    sf = _c(scale_factors, np.float32)
    valid, u, v = _c(valid, np.uint8), _c(u, np.float32), _c(v, np.float32)
    level, qd = _c(level, np.int32), _c(qdesc, np.uint8)
    km = np.full(len(kps), -1, np.int32)
    n = _L().orc_search_sim3(len(kps), _p(kps), _p(desc), _p(b), _p(m), _p(sf), len(valid), _p(valid), _p(u), _p(v),
                             _p(level), _p(qd), int(th), float(ratio_hamming), _p(km))
    return n, km


def search_initialization(kps1, desc1, kps2, desc2, bounds, prev_matched, window, nnratio, check_ori=True):
    """M5.  Returns (nmatches, matches12[n1], prevMatched_after[n1,2])."""
    kps1, desc1, kps2, desc2, b = _kp(kps1), _c(desc1, np.uint8), _kp(kps2), _c(desc2, np.uint8), _bounds(bounds)
    pm = _c(prev_matched, np.float32).copy()
    m12 = np.full(len(kps1), -1, np.int32)
    n = _L().orc_search_initialization(len(kps1), _p(kps1), _p(desc1), len(kps2), _p(kps2), _p(desc2), _p(b), _p(pm),
                                       int(window), float(nnratio), int(check_ori), _p(m12))
    return n, m12, pm


def _fv(fv):
    node, off, idx = fv
    return _c(node, np.int32), _c(off, np.int32), _c(idx, np.int32)


def search_bow_kf_frame(kps1, desc1, fv1, elig1, kps2, desc2, fv2, nnratio, check_ori=True):
    """M4a.  Returns (nmatches, matchF[n2])."""
    kps1, desc1, kps2, desc2 = _kp(kps1), _c(desc1, np.uint8), _kp(kps2), _c(desc2, np.uint8)
    n1, o1, i1 = _fv(fv1); n2, o2, i2 = _fv(fv2)
    e1 = _c(elig1, np.uint8)
    out = np.full(len(kps2), -1, np.int32)
    n = _L().orc_search_bow_kf_frame(len(kps1), _p(kps1), _p(desc1), len(n1), _p(n1), _p(o1), _p(i1), _p(e1), len(kps2),
                                     _p(kps2), _p(desc2), len(n2), _p(n2), _p(o2), _p(i2), float(nnratio), int(check_ori), _p(out))
    return n, out


def search_bow_kf_kf(kps1, desc1, fv1, elig1, kps2, desc2, fv2, elig2, nnratio, check_ori=True):
    """M4b.  Returns (nmatches, matches12[n1])."""
    kps1, desc1, kps2, desc2 = _kp(kps1), _c(desc1, np.uint8), _kp(kps2), _c(desc2, np.uint8)
    n1, o1, i1 = _fv(fv1); n2, o2, i2 = _fv(fv2)
    e1, e2 = _c(elig1, np.uint8), _c(elig2, np.uint8)
    out = np.full(len(kps1), -1, np.int32)
    n = _L().orc_search_bow_kf_kf(len(kps1), _p(kps1), _p(desc1), len(n1), _p(n1), _p(o1), _p(i1), _p(e1), len(kps2),
                                  _p(kps2), _p(desc2), len(n2), _p(n2), _p(o2), _p(i2), _p(e2), float(nnratio), int(check_ori), _p(out))
    return n, out


def search_triangulation(kps1, desc1, fv1, free1, kps2, desc2, fv2, free2, ep, scale_factors2, F12, sigma2, coarse=False,
                         check_ori=False):
    """M6.  Returns (nmatches, matches12[n1])."""
    kps1, desc1, kps2, desc2 = _kp(kps1), _c(desc1, np.uint8), _kp(kps2), _c(desc2, np.uint8)
    n1, o1, i1 = _fv(fv1); n2, o2, i2 = _fv(fv2)
    f1, f2 = _c(free1, np.uint8), _c(free2, np.uint8)
    ep, sf, F, s2 = _c(ep, np.float32), _c(scale_factors2, np.float32), _c(F12, np.float32).reshape(9), _c(sigma2, np.float32)
    out = np.full(len(kps1), -1, np.int32)
    n = _L().orc_search_triangulation(len(kps1), _p(kps1), _p(desc1), len(n1), _p(n1), _p(o1), _p(i1), _p(f1), len(kps2),
                                      _p(kps2), _p(desc2), len(n2), _p(n2), _p(o2), _p(i2), _p(f2), _p(ep), _p(sf), _p(F),
                                      _p(s2), int(coarse), int(check_ori), _p(out))
    return n, out


def fuse_candidates(kps, desc, bounds, scale_factors, inv_sigma2, valid, u, v, level, qdesc, th, chi2_gate=True):
    """M9 candidate part.  Returns (bestIdx[nq], bestDist[nq])."""
    kps, desc, b = _kp(kps), _c(desc, np.uint8), _bounds(bounds)
    sf, is2 = _c(scale_factors, np.float32), _c(inv_sigma2, np.float32)
    valid, u, v = _c(valid, np.uint8), _c(u, np.float32), _c(v, np.float32)
    level, qd = _c(level, np.int32), _c(qdesc, np.uint8)
    bi = np.empty(len(valid), np.int32); bd = np.empty(len(valid), np.int32)
    _L().orc_fuse_candidates(len(kps), _p(kps), _p(desc), _p(b), _p(sf), _p(is2), len(valid), _p(valid), _p(u), _p(v),
                             _p(level), _p(qd), float(th), int(chi2_gate), _p(bi), _p(bd))
    return bi, bd


def search_by_sim3(kps1, desc1, bounds1, kps2, desc2, bounds2, sf1, sf2, valid1, u12, v12, level12, qdesc1, valid2, u21, v21,
                   level21, qdesc2, th):
    """M8 SearchBySim3.  Returns (nFound, matches12[n1])."""
    kps1, desc1, b1 = _kp(kps1), _c(desc1, np.uint8), _bounds(bounds1)
    kps2, desc2, b2 = _kp(kps2), _c(desc2, np.uint8), _bounds(bounds2)
    sf1, sf2 = _c(sf1, np.float32), _c(sf2, np.float32)
    a = [_c(valid1, np.uint8), _c(u12, np.float32), _c(v12, np.float32), _c(level12, np.int32), _c(qdesc1, np.uint8)]
    b = [_c(valid2, np.uint8), _c(u21, np.float32), _c(v21, np.float32), _c(level21, np.int32), _c(qdesc2, np.uint8)]
    out = np.full(len(kps1), -1, np.int32)
    n = _L().orc_search_by_sim3(len(kps1), _p(kps1), _p(desc1), _p(b1), len(kps2), _p(kps2), _p(desc2), _p(b2), _p(sf1), _p(sf2),
                                *[_p(x) for x in a], *[_p(x) for x in b], float(th), _p(out))
    return n, out


def stereo_matches(kpsL, descL, kpsR, descR, pyrL, pyrR, scale, inv_scale, mb, mbf):
    """Frame::ComputeStereoMatches.  pyrL / pyrR: lists of 2-D uint8 level planes.  Returns (kept, mvuRight[N], mvDepth[N])."""
    kpsL, descL, kpsR, descR = _kp(kpsL), _c(descL, np.uint8), _kp(kpsR), _c(descR, np.uint8)
    pl = [np.ascontiguousarray(a, np.uint8) for a in pyrL]
    pr = [np.ascontiguousarray(a, np.uint8) for a in pyrR]
    n = len(pl)
    ptr = lambda planes: (C.c_void_p * n)(*[a.ctypes.data for a in planes])
    sl = np.array([a.shape[1] for a in pl], np.int32)
    sr = np.array([a.shape[1] for a in pr], np.int32)
    sf, isf = _c(scale, np.float32), _c(inv_scale, np.float32)
    ur = np.empty(len(kpsL), np.float32); dp = np.empty(len(kpsL), np.float32)
    kept = _L().orc_stereo_matches(len(kpsL), _p(kpsL), _p(descL), len(kpsR), _p(kpsR), _p(descR), n, ptr(pl), _p(sl), ptr(pr),
                                   _p(sr), _p(sr), _p(sf), _p(isf), float(mb), float(mbf), _p(ur), _p(dp))
    return kept, ur, dp


def knn(q, db, k, threads=1):
    """M11.  Returns (dist[nq,k] uint16, idx[nq,k] int32); ties -> lowest row."""
    q, db = _c(q, np.uint8), _c(db, np.uint8)
    dist = np.empty((len(q), k), np.uint16); idx = np.empty((len(q), k), np.int32)
    _L().orc_knn(_p(q), len(q), _p(db), len(db), int(k), _p(dist), _p(idx), int(threads))
    return dist, idx


def distinctive_descriptor(desc):
    desc = _c(desc, np.uint8)
    return int(_L().orc_distinctive_descriptor(_p(desc), len(desc)))


def undistort_points(xy, K, dist):
    """cv::undistortPoints(xy, K, dist, R=I, P=K) as Frame::UndistortKeyPoints calls it (src/Frame.cc:768).
    K = (fx, fy, cx, cy)."""
    xy, K, dist = _c(xy, np.float32).reshape(-1, 2), _c(K, np.float32), _c(dist, np.float32)
    out = np.empty_like(xy)
    _L().orc_undistort_points(_p(xy), len(xy), _p(K), _p(dist), len(dist), _p(out))
    return out


def bow_transform(voc, desc, levelsup=4):
    """DBoW2 transform(features, BowVector, FeatureVector, levelsup).  voc: dict with L, child_off, children, desc, word_id,
    weight, weighting, norm (see orc_matcher.h VocabView).  Returns dict(word, node, bow_word, bow_value, fv_node, fv_off,
    fv_idx)."""
    desc = _c(desc, np.uint8).reshape(-1, 32)
    n = len(desc)
    co, ch = _c(voc["child_off"], np.int32), _c(voc["children"], np.int32)
    nd, wi, we = _c(voc["desc"], np.uint8), _c(voc["word_id"], np.int32), _c(voc["weight"], np.float64)
    fw, fn = np.empty(n, np.int32), np.empty(n, np.int32)
    bw, bv = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.float64)
    vn, vo, vi = np.empty(max(n, 1), np.int32), np.empty(n + 1, np.int32), np.empty(max(n, 1), np.int32)
    nb, nf = C.c_int(0), C.c_int(0)
    _L().orc_bow_transform(len(co) - 1, int(voc["L"]), _p(co), _p(ch), _p(nd), _p(wi), _p(we), int(voc["weighting"]),
                           int(voc["norm"]), _p(desc), n, int(levelsup), _p(fw), _p(fn), C.byref(nb), _p(bw), _p(bv),
                           C.byref(nf), _p(vn), _p(vo), _p(vi))
    nb, nf = nb.value, nf.value
    return dict(word=fw, node=fn, bow_word=bw[:nb].copy(), bow_value=bv[:nb].copy(), fv_node=vn[:nf].copy(),
                fv_off=vo[:nf + 1].copy(), fv_idx=vi[:vo[nf]].copy())
