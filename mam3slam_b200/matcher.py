"""Host-side mirror of the reference's ORBmatcher (include/ORBmatcher.h:36-103) over the C ABI (include/orbm.h).

The reference's methods take Frame / KeyFrame / MapPoint pointer graphs; here a `MatchFrame` is the device mirror of
what the searches read from a Frame or KeyFrame (mvKeysUn, mDescriptors, image bounds, the 64x48 grid) and every
MapPoint-dependent predicate arrives as a per-element array, exactly as the C++ adapter passes it.  Method names follow
the reference; all Hamming work runs in the CUDA library (no CPU fallback).
"""
import ctypes as C
import threading
import numpy as np

from ._lib import lib, check_m, OrbmFrameView, OrbmFeatVec, OrbmVocabView
from .extractor import KEYPOINT_DTYPE


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dt)


class MatchFrame:
    """Device mirror of a Frame / KeyFrame for the searches (orbm_frame)."""

    def __init__(self, matcher, kps, desc, bounds, on_device=False, n=None):
        self._m = matcher
        self._L = matcher._L
        self._f = C.c_void_p()
        self._keep = None
        view = self._view(kps, desc, bounds, on_device, n)
        check_m(self._L.orbm_frame_create(matcher._h, C.byref(view), C.byref(self._f)), matcher._h, "orbm_frame_create")

    def _view(self, kps, desc, bounds, on_device, n):
        b = [float(x) for x in bounds]
        if on_device:
            self.n = int(n)
            return OrbmFrameView(self.n, C.c_void_p(kps), C.c_void_p(desc), b[0], b[1], b[2], b[3], 1)
        kps = np.ascontiguousarray(kps)
        assert kps.dtype.itemsize == 28, "cv::KeyPoint layout expected"
        desc = _c(desc, np.uint8)
        self.n = len(kps)
        self._keep = (kps, desc)
        return OrbmFrameView(self.n, kps.ctypes.data, desc.ctypes.data, b[0], b[1], b[2], b[3], 0)

    def update(self, kps, desc, bounds, on_device=False, n=None):
        view = self._view(kps, desc, bounds, on_device, n)
        check_m(self._L.orbm_frame_update(self._m._h, self._f, C.byref(view)), self._m._h, "orbm_frame_update")

    def grid(self):
        """(off[64*48+1], idx[...]) — Frame::AssignFeaturesToGrid as CSR, cell = ix*48+iy."""
        off = np.empty(64 * 48 + 1, np.int32)
        idx = np.empty(max(self.n, 1), np.int32)
        check_m(self._L.orbm_frame_grid(self._m._h, self._f, _p(off), _p(idx)), self._m._h, "orbm_frame_grid")
        return off, idx[:off[-1]].copy()

    def GetFeaturesInArea(self, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.empty(max(self.n, 1), np.int32)
        n = C.c_int(0)
        check_m(self._L.orbm_features_in_area(self._m._h, self._f, float(x), float(y), float(r), int(minLevel), int(maxLevel),
                                              _p(out), len(out), C.byref(n)), self._m._h, "orbm_features_in_area")
        return out[:n.value].copy()

    def close(self):
        if getattr(self, "_f", None) and self._m._h:
            self._L.orbm_frame_destroy(self._m._h, self._f)
            self._f = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _fv(fv, keep):
    node, off, idx = (_c(a, np.int32) for a in fv)
    keep.extend([node, off, idx])
    return OrbmFeatVec(len(node), node.ctypes.data, off.ctypes.data, idx.ctypes.data)


class KnnDatabase:
    """Descriptor rows resident in HBM for brute-force k-NN (orbm_db); row_offset = global id of the first row."""

    def __init__(self, matcher, rows, row_offset=0, device_ptr=None, n=None):
        self._m = matcher
        self._db = C.c_void_p()
        if device_ptr is not None:
            self.n = int(n)
            check_m(matcher._L.orbm_db_create(matcher._h, C.c_void_p(device_ptr), self.n, 1, int(row_offset), C.byref(self._db)),
                    matcher._h, "orbm_db_create")
        else:
            rows = _c(rows, np.uint8)
            self.n = len(rows)
            check_m(matcher._L.orbm_db_create(matcher._h, _p(rows), self.n, 0, int(row_offset), C.byref(self._db)),
                    matcher._h, "orbm_db_create")

    def close(self):
        if getattr(self, "_db", None) and self._m._h:
            self._m._L.orbm_db_destroy(self._m._h, self._db)
            self._db = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class KnnExchange:
    """This rank's side of the sharded k-NN over peer memory (orbm_xchg, include/orbm.h): one buffer per rank, every rank
    maps every other rank's buffer, and ORBmatcher.knn_sharded does scan + all-gather + merge inside its two kernels.

    `connect(all_gather)` does the setup for one process per GPU: `all_gather(bytes) -> [bytes] * world` is whatever the host
    has - with torch.distributed: `lambda b: (lambda out: (dist.all_gather_object(out, b), out)[1])([None] * world)`.
    `link(peers)` is the single-process variant (a list of the KnnExchange objects of all ranks)."""
    HANDLE_BYTES = 64

    def __init__(self, matcher, rank, world, max_nq, max_k):
        self._m, self.rank, self.world = matcher, int(rank), int(world)
        self._x = C.c_void_p()
        check_m(matcher._L.orbm_xchg_create(matcher._h, self.rank, self.world, int(max_nq), int(max_k), C.byref(self._x)),
                matcher._h, "orbm_xchg_create")

    def ipc_handle(self):
        buf = C.create_string_buffer(self.HANDLE_BYTES)
        check_m(self._m._L.orbm_xchg_ipc_handle(self._x, buf), self._m._h, "orbm_xchg_ipc_handle")
        return buf.raw

    def connect(self, all_gather):
        handles = all_gather(self.ipc_handle())
        assert len(handles) == self.world
        for r, h in enumerate(handles):
            if r != self.rank:
                check_m(self._m._L.orbm_xchg_open(self._x, r, C.create_string_buffer(bytes(h), self.HANDLE_BYTES)), self._m._h,
                        "orbm_xchg_open")
        return self

    def link(self, peers):
        for r, p in enumerate(peers):
            if r != self.rank:
                check_m(self._m._L.orbm_xchg_open_local(self._x, r, p._x), self._m._h, "orbm_xchg_open_local")
        return self

    def status(self):
        return int(self._m._L.orbm_xchg_status(self._x))

    def close(self):
        if getattr(self, "_x", None) and self._m._h:
            self._m._L.orbm_xchg_destroy(self._x)
            self._x = None


class ORBVocabulary:
    """DBoW2 vocabulary tree resident in HBM (orbm_vocab) — the part of ORBVocabulary (include/ORBVocabulary.h;
    Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) that Frame::ComputeBoW / KeyFrame::ComputeBoW use.  The tree arrives as
    plain arrays (see include/orbm.h orbm_vocab_view); TF_IDF weighting and L1 norm are ORBvoc's settings."""
    TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3

    def __init__(self, matcher, levels, child_off, children, node_desc, word_id, weight, weighting=0, norm=1):
        self._m = matcher
        self._keep = [_c(child_off, np.int32), _c(children, np.int32), _c(node_desc, np.uint8), _c(word_id, np.int32),
                      _c(weight, np.float64)]
        co, ch, nd, wi, we = self._keep
        view = OrbmVocabView(len(co) - 1, int(levels), co.ctypes.data, ch.ctypes.data, nd.ctypes.data, wi.ctypes.data,
                             we.ctypes.data, int(weighting), int(norm))
        self._v = C.c_void_p()
        check_m(matcher._L.orbm_vocab_create(matcher._h, C.byref(view), C.byref(self._v)), matcher._h, "orbm_vocab_create")

    def transform(self, desc, levelsup=4):
        """transform(features, BowVector&, FeatureVector&, levelsup): returns ((bow_word, bow_value), (fv_node, fv_off,
        fv_idx), feat_word, feat_node)."""
        desc = _c(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        fw, fn = np.empty(n, np.int32), np.empty(n, np.int32)
        bw, bv = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.float64)
        vn, vo, vi = np.empty(max(n, 1), np.int32), np.empty(n + 1, np.int32), np.empty(max(n, 1), np.int32)
        nb, nf = C.c_int(0), C.c_int(0)
        m = self._m
        check_m(m._L.orbm_bow_transform(m._h, self._v, _p(desc), n, int(levelsup), _p(fw), _p(fn), C.byref(nb), _p(bw), _p(bv),
                                        C.byref(nf), _p(vn), _p(vo), _p(vi)), m._h, "orbm_bow_transform")
        nb, nf = nb.value, nf.value
        return (bw[:nb].copy(), bv[:nb].copy()), (vn[:nf].copy(), vo[:nf + 1].copy(), vi[:vo[nf]].copy()), fw, fn

    def transform_device(self, nframes, d_desc, d_counts, frame_stride, levelsup, d_feat_word, d_feat_node, count_stride=1,
                         stream=None):
        m = self._m
        check_m(m._L.orbm_bow_transform_device(m._h, self._v, int(nframes), C.c_void_p(d_desc), C.c_void_p(d_counts),
                                               int(frame_stride), int(count_stride), int(levelsup), C.c_void_p(d_feat_word),
                                               C.c_void_p(d_feat_node), C.c_void_p(stream or 0)), m._h, "orbm_bow_transform_device")

    def close(self):
        if getattr(self, "_v", None) and self._m._h:
            self._m._L.orbm_vocab_destroy(self._m._h, self._v)
            self._v = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ORBmatcher:
    TH_HIGH = 100
    TH_LOW = 50
    HISTO_LENGTH = 30

    # The reference constructs an ORBmatcher per call (a stack local in every tracking / mapping / loop-closing function), so
    # construction must be free: like the C++ adapter (cpp/ORBmatcher.h), the CUDA-side state - orbm handle (stream +
    # scratch) and two device frame slots - comes from a per-thread pool.  __init__ pops a set for the device, close() pushes it
    # back; drain_pool() destroys what the calling thread's pool holds.
    _POOL = threading.local()

    @classmethod
    def _pool(cls):
        if not hasattr(cls._POOL, "sets"):
            cls._POOL.sets = []
        return cls._POOL.sets

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self._L = lib()
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        self._device = int(device)
        sets = self._pool()
        for i in range(len(sets) - 1, -1, -1):
            if sets[i][0] == self._device:
                _, self._h, self._slots = sets.pop(i)
                return
        self._h = C.c_void_p()
        self._slots = [None, None]
        check_m(self._L.orbm_create(self._device, C.byref(self._h)), None, "orbm_create")

    def close(self):
        if getattr(self, "_h", None):
            self._pool().append((self._device, self._h, self._slots))
            self._h = None
            self._slots = [None, None]

    def slot(self, i, kps, desc, bounds, on_device=False, n=None):
        """One of the matcher's two persistent device frame slots (the adapter's a_ / b_), bound to a frame: created on first
        use, afterwards only re-uploaded - no allocation in steady state, also across pooled ORBmatcher objects."""
        F = self._slots[i]
        if F is None:
            F = self._slots[i] = MatchFrame(self, kps, desc, bounds, on_device, n)
        else:
            F._m = self
            F.update(kps, desc, bounds, on_device, n)
        return F

    @classmethod
    def drain_pool(cls):
        L = lib()
        for _, h, slots in cls._pool():
            for F in slots:
                if F is not None and F._f:
                    L.orbm_frame_destroy(h, F._f)
                    F._f = None
            L.orbm_destroy(h)
        cls._pool().clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def launch_count(self):
        return int(self._L.orbm_launch_count(self._h))

    def last_evaluations(self):
        return int(self._L.orbm_last_evaluations(self._h))

    def stream(self):
        return self._L.orbm_stream(self._h)

    # ---- M0 ----
    @staticmethod
    def DescriptorDistance(a, b):
        a, b = _c(a, np.uint8), _c(b, np.uint8)
        return int(lib().orbm_distance(_p(a), _p(b)))

    def frame(self, kps, desc, bounds, on_device=False, n=None):
        return MatchFrame(self, kps, desc, bounds, on_device, n)

    # ---- M1: SearchByProjection(Frame&, vector<MapPoint*>&, th, ...) ----
    def SearchByProjection_LocalMap(self, F, occupied, scale_factors, valid, proj_x, proj_y, level, view_cos, qdesc, th=1.0,
                                    q_occupies=None):
        """q_occupies[q] = the query's MapPoint has Observations() > 0 (src/ORBmatcher.cc:92-94); None = every match occupies."""
        qo = None if q_occupies is None else _c(q_occupies, np.uint8)
        occ = _c(occupied, np.uint8).copy()
        sf = _c(scale_factors, np.float32)
        valid, px, py = _c(valid, np.uint8), _c(proj_x, np.float32), _c(proj_y, np.float32)
        level, vc, qd = _c(level, np.int32), _c(view_cos, np.float32), _c(qdesc, np.uint8)
        km = np.full(F.n, -1, np.int32)
        nm = C.c_int(0)
        check_m(self._L.orbm_search_local_map(self._h, F._f, _p(occ), _p(sf), len(sf), len(valid), _p(valid), _p(px), _p(py),
                                              _p(level), _p(vc), _p(qd), float(th), self.mfNNratio, _p(km), C.byref(nm),
                                              None if qo is None else _p(qo)),
                self._h, "orbm_search_local_map")
        return nm.value, km, occ

    # ---- M2 / M3: SearchByProjection(Frame& Cur, Frame& Last, th, bMono) / (Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) ----
    def SearchByProjection_Frame(self, F, occupied, scale_factors, valid, u, v, octave, qangle, qdesc, th, ORBdist=100,
                                 q_occupies=None):
        qo = None if q_occupies is None else _c(q_occupies, np.uint8)          # src/ORBmatcher.cc:1745-1747
        occ = _c(occupied, np.uint8).copy()
        sf = _c(scale_factors, np.float32)
        valid, u, v = _c(valid, np.uint8), _c(u, np.float32), _c(v, np.float32)
        octave, qa, qd = _c(octave, np.int32), _c(qangle, np.float32), _c(qdesc, np.uint8)
        km = np.full(F.n, -1, np.int32)
        nm = C.c_int(0)
        check_m(self._L.orbm_search_frame(self._h, F._f, _p(occ), _p(sf), len(sf), len(valid), _p(valid), _p(u), _p(v), _p(octave),
                                          _p(qa), _p(qd), float(th), int(ORBdist), int(self.mbCheckOrientation), _p(km),
                                          C.byref(nm), None if qo is None else _p(qo)), self._h, "orbm_search_frame")
        return nm.value, km

    def track_batch_device(self, nframes, d_kps, d_desc, d_counts, frame_stride, bounds, scale_factors, th, d_kp_match,
                           d_nmatches, ORBdist=100, count_stride=2, stream=None):
        """Device-resident SearchByProjection(Cur, Last) over consecutive frames of an extractor batch (orbm_track_batch_device)."""
        b, sf = _c(bounds, np.float32), _c(scale_factors, np.float32)
        check_m(self._L.orbm_track_batch_device(self._h, int(nframes), C.c_void_p(d_kps), C.c_void_p(d_desc), C.c_void_p(d_counts),
                                                int(frame_stride), int(count_stride), _p(b), _p(sf), len(sf), float(th), int(ORBdist),
                                                int(self.mbCheckOrientation), C.c_void_p(d_kp_match), C.c_void_p(d_nmatches),
                                                C.c_void_p(stream or 0)), self._h, "orbm_track_batch_device")

    def track_evaluations(self):
        return int(self._L.orbm_track_evaluations(self._h))

    # ---- M7: SearchByProjection(KeyFrame*, Sim3, points, matched, th, ratioHamming) ----
    def SearchByProjection_Sim3(self, KF, matched, scale_factors, valid, u, v, level, qdesc, th, ratioHamming=1.0):
        m = _c(matched, np.uint8).copy()
        sf = _c(scale_factors, np.float32)
        valid, u, v = _c(valid, np.uint8), _c(u, np.float32), _c(v, np.float32)
        level, qd = _c(level, np.int32), _c(qdesc, np.uint8)
        km = np.full(KF.n, -1, np.int32)
        nm = C.c_int(0)
        check_m(self._L.orbm_search_sim3(self._h, KF._f, _p(m), _p(sf), len(sf), len(valid), _p(valid), _p(u), _p(v), _p(level),
                                         _p(qd), int(th), float(ratioHamming), _p(km), C.byref(nm)), self._h, "orbm_search_sim3")
        return nm.value, km

    # ---- M5: SearchForInitialization ----
    def SearchForInitialization(self, F1, F2, vbPrevMatched, windowSize=10):
        pm = _c(vbPrevMatched, np.float32).copy()
        m12 = np.full(F1.n, -1, np.int32)
        nm = C.c_int(0)
        check_m(self._L.orbm_search_initialization(self._h, F1._f, F2._f, _p(pm), int(windowSize), self.mfNNratio,
                                                   int(self.mbCheckOrientation), _p(m12), C.byref(nm)), self._h,
                "orbm_search_initialization")
        return nm.value, m12, pm

    # ---- M4: SearchByBoW ----
    def SearchByBoW_KF_Frame(self, KF, fvKF, elig1, F, fvF):
        keep = []
        a, b = _fv(fvKF, keep), _fv(fvF, keep)
        e1 = _c(elig1, np.uint8)
        out = np.full(F.n, -1, np.int32)
        nm = C.c_int(0)
        check_m(self._L.orbm_search_bow_kf_frame(self._h, KF._f, C.byref(a), _p(e1), F._f, C.byref(b), self.mfNNratio,
                                                 int(self.mbCheckOrientation), _p(out), C.byref(nm)), self._h,
                "orbm_search_bow_kf_frame")
        return nm.value, out

    def SearchByBoW_KF_KF(self, KF1, fv1, elig1, KF2, fv2, elig2):
        keep = []
        a, b = _fv(fv1, keep), _fv(fv2, keep)
        e1, e2 = _c(elig1, np.uint8), _c(elig2, np.uint8)
        out = np.full(KF1.n, -1, np.int32)
        nm = C.c_int(0)
        check_m(self._L.orbm_search_bow_kf_kf(self._h, KF1._f, C.byref(a), _p(e1), KF2._f, C.byref(b), _p(e2), self.mfNNratio,
                                              int(self.mbCheckOrientation), _p(out), C.byref(nm)), self._h,
                "orbm_search_bow_kf_kf")
        return nm.value, out

    # ---- M6: SearchForTriangulation ----
    def SearchForTriangulation(self, KF1, fv1, free1, KF2, fv2, free2, ep, scale_factors2, F12, sigma2, bCoarse=False):
        keep = []
        a, b = _fv(fv1, keep), _fv(fv2, keep)
        f1, f2 = _c(free1, np.uint8), _c(free2, np.uint8)
        ep, sf = _c(ep, np.float32), _c(scale_factors2, np.float32)
        F, s2 = _c(F12, np.float32).reshape(9), _c(sigma2, np.float32)
        out = np.full(KF1.n, -1, np.int32)
        nm = C.c_int(0)
        check_m(self._L.orbm_search_triangulation(self._h, KF1._f, C.byref(a), _p(f1), KF2._f, C.byref(b), _p(f2), _p(ep), _p(sf),
                                                  _p(F), _p(s2), len(sf), int(bCoarse), int(self.mbCheckOrientation), _p(out),
                                                  C.byref(nm)), self._h, "orbm_search_triangulation")
        return nm.value, out

    # ---- M9: candidate part of Fuse ----
    def Fuse_candidates(self, KF, scale_factors, inv_sigma2, valid, u, v, level, qdesc, th, chi2_gate=True):
        sf, is2 = _c(scale_factors, np.float32), _c(inv_sigma2, np.float32)
        valid, u, v = _c(valid, np.uint8), _c(u, np.float32), _c(v, np.float32)
        level, qd = _c(level, np.int32), _c(qdesc, np.uint8)
        bi = np.empty(len(valid), np.int32)
        bd = np.empty(len(valid), np.int32)
        check_m(self._L.orbm_fuse_candidates(self._h, KF._f, _p(sf), _p(is2), len(sf), len(valid), _p(valid), _p(u), _p(v),
                                             _p(level), _p(qd), float(th), int(chi2_gate), _p(bi), _p(bd)), self._h,
                "orbm_fuse_candidates")
        return bi, bd

    # ---- M8: SearchBySim3 (src/ORBmatcher.cc:1457-1674) ----
    def SearchBySim3(self, KF1, KF2, scale_factors1, scale_factors2, valid1, u12, v12, level12, qdesc1, valid2, u21, v21,
                     level21, qdesc2, th):
        sf1, sf2 = _c(scale_factors1, np.float32), _c(scale_factors2, np.float32)
        a = [_c(valid1, np.uint8), _c(u12, np.float32), _c(v12, np.float32), _c(level12, np.int32), _c(qdesc1, np.uint8)]
        b = [_c(valid2, np.uint8), _c(u21, np.float32), _c(v21, np.float32), _c(level21, np.int32), _c(qdesc2, np.uint8)]
        if len(a[0]) != KF1.n or len(b[0]) != KF2.n:
            raise ValueError("one query per keypoint slot of each keyframe")
        out = np.full(KF1.n, -1, np.int32)
        nf = C.c_int(0)
        check_m(self._L.orbm_search_by_sim3(self._h, KF1._f, KF2._f, _p(sf1), _p(sf2), len(sf1), *[_p(x) for x in a],
                                            *[_p(x) for x in b], float(th), _p(out), C.byref(nf)), self._h,
                "orbm_search_by_sim3")
        return nf.value, out

    # ---- M11: brute-force k-NN ----
    def database(self, rows, row_offset=0):
        return KnnDatabase(self, rows, row_offset)

    def knn(self, q, db, k):
        """(dist[nq,k] uint16, idx[nq,k] int32); `db` = KnnDatabase or an array of rows (uploaded for this call)."""
        own = None
        if not isinstance(db, KnnDatabase):
            own = db = KnnDatabase(self, db)
        q = _c(q, np.uint8)
        dist = np.empty((len(q), k), np.uint16)
        idx = np.empty((len(q), k), np.int32)
        check_m(self._L.orbm_knn(self._h, db._db, _p(q), len(q), int(k), _p(dist), _p(idx)), self._h, "orbm_knn")
        if own is not None:
            own.close()
        return dist, idx

    def knn_device(self, db, d_q, nq, k, d_dist, d_idx, stream=None):
        check_m(self._L.orbm_knn_device(self._h, db._db, C.c_void_p(d_q), int(nq), int(k), C.c_void_p(d_dist), C.c_void_p(d_idx),
                                        C.c_void_p(stream or 0)), self._h, "orbm_knn_device")

    def knn_device_tc(self, db, d_q, nq, k, d_dist, d_idx, stream=None):
        """orbm_knn_device_tc: the brute-force search with the dot products on the tensor cores (tcgen05.mma kind::i8); same
        arguments and the same result as knn_device (include/orbm.h)."""
        check_m(self._L.orbm_knn_device_tc(self._h, db._db, C.c_void_p(d_q), int(nq), int(k), C.c_void_p(d_dist), C.c_void_p(d_idx),
                                           C.c_void_p(stream or 0)), self._h, "orbm_knn_device_tc")

    def knn_device_static(self, db, d_q, nq, k, d_dist, d_idx, stream=None):
        """orbm_knn_device_static: the same search with its scratch in a fixed-size static device array; the call allocates
        nothing (include/orbm.h).  No counterpart in the reference."""
        check_m(self._L.orbm_knn_device_static(self._h, db._db, C.c_void_p(d_q), int(nq), int(k), C.c_void_p(d_dist),
                                               C.c_void_p(d_idx), C.c_void_p(stream or 0)), self._h, "orbm_knn_device_static")

    def knn_sharded(self, xchg, db, d_q, nq, k, d_dist, d_idx, stream=None):
        """orbm_knn_sharded: this rank's shard scan, the exchange of the per-rank lists over NVLink peer stores and the final
        merge, all asynchronous on `stream`; every rank calls it with the same queries (include/orbm.h)."""
        check_m(self._L.orbm_knn_sharded(self._h, xchg._x, db._db, C.c_void_p(d_q), int(nq), int(k), C.c_void_p(d_dist),
                                         C.c_void_p(d_idx), C.c_void_p(stream or 0)), self._h, "orbm_knn_sharded")

    def knn_merge_device(self, d_dist, d_idx, shard_stride_bytes, nshards, nq, k, d_out_dist, d_out_idx, stream=None):
        """Device-side merge of all-gathered per-shard top-k lists (include/orbm.h); pointers are device addresses."""
        check_m(self._L.orbm_knn_merge_device(self._h, C.c_void_p(d_dist), C.c_void_p(d_idx), int(shard_stride_bytes), int(nshards),
                                              int(nq), int(k), C.c_void_p(d_out_dist), C.c_void_p(d_out_idx),
                                              C.c_void_p(stream or 0)), self._h, "orbm_knn_merge_device")

    @staticmethod
    def knn_merge(dist, idx):
        """dist/idx: [nshards, nq, k] per-shard lists with global row ids -> merged [nq, k]."""
        dist, idx = _c(dist, np.uint16), _c(idx, np.int32)
        s, nq, k = dist.shape
        od = np.empty((nq, k), np.uint16)
        oi = np.empty((nq, k), np.int32)
        st = lib().orbm_knn_merge(_p(dist), _p(idx), s, nq, k, _p(od), _p(oi))
        check_m(st, None, "orbm_knn_merge")
        return od, oi

    # ---- 8(f): MapPoint::ComputeDistinctiveDescriptors selection ----
    def ComputeDistinctiveDescriptors(self, desc, off):
        desc, off = _c(desc, np.uint8), _c(off, np.int32)
        best = np.empty(len(off) - 1, np.int32)
        check_m(self._L.orbm_distinctive_descriptors(self._h, _p(desc), _p(off), len(off) - 1, _p(best)), self._h,
                "orbm_distinctive_descriptors")
        return best

    # ---- 8(f): Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:747-812) ----
    def UndistortPoints(self, xy, K4, dist):
        """cv::undistortPoints(xy, K, dist, R=I, P=K); K4 = (fx, fy, cx, cy).  As in the reference, callers skip the call when
        dist[0] == 0."""
        xy = _c(xy, np.float32).reshape(-1, 2)
        K4, dist = _c(K4, np.float32), _c(dist, np.float32)
        out = np.empty_like(xy)
        check_m(self._L.orbm_undistort_points(self._h, _p(xy), len(xy), _p(K4), _p(dist), len(dist), _p(out)), self._h,
                "orbm_undistort_points")
        return out

    def undistort_keypoints_device(self, nframes, d_kps, d_counts, frame_stride, K4, dist, d_kps_un, count_stride=1, stream=None):
        K4, dist = _c(K4, np.float32), _c(dist, np.float32)
        check_m(self._L.orbm_undistort_keypoints_device(self._h, int(nframes), C.c_void_p(d_kps), C.c_void_p(d_counts),
                                                        int(frame_stride), int(count_stride), _p(K4), _p(dist), len(dist),
                                                        C.c_void_p(d_kps_un), C.c_void_p(stream or 0)), self._h,
                "orbm_undistort_keypoints_device")

    def vocabulary(self, levels, child_off, children, node_desc, word_id, weight, weighting=0, norm=1):
        return ORBVocabulary(self, levels, child_off, children, node_desc, word_id, weight, weighting, norm)
