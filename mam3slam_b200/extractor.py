"""Host-side mirror of the reference's ORBextractor (include/ORBextractor.h:43-109) over the C ABI (include/orbx.h).

Same constructor arguments, same call semantics and getters as the C++ class, so the parity tests read like calls
into the reference:  mono = extractor(image, None, keypoints_out, descriptors_out, vLappingArea)  becomes
mono, keypoints, descriptors = extractor(image, None, vLappingArea).  All compute runs in the CUDA library.
"""
import ctypes as C
import numpy as np

from ._lib import lib, check, OrbxConfig

# numpy view of orbx_keypoint == cv::KeyPoint's binary layout
KEYPOINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                           ("octave", "<i4"), ("class_id", "<i4")])
assert KEYPOINT_DTYPE.itemsize == 28


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class ORBextractor:
    HARRIS_SCORE = 0
    FAST_SCORE = 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0, max_batch=1):
        self._L = lib()
        self.nfeatures, self.scaleFactor, self.nlevels = int(nfeatures), float(scaleFactor), int(nlevels)
        self.iniThFAST, self.minThFAST = int(iniThFAST), int(minThFAST)
        self.max_batch = int(max_batch)
        cfg = OrbxConfig(self.nfeatures, self.scaleFactor, self.nlevels, self.iniThFAST, self.minThFAST, int(device),
                         self.max_batch)
        h = C.c_void_p()
        check(self._L.orbx_create(C.byref(cfg), C.byref(h)), None, "orbx_create")
        self._h = h
        n = self.nlevels
        self._tables = [np.empty(n, np.float32) for _ in range(4)] + [np.empty(n, np.int32)]
        check(self._L.orbx_get_tables(h, *[_p(t) for t in self._tables]), h, "orbx_get_tables")
        self._frames = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- getters (include/ORBextractor.h:61-81) ----
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self.scaleFactor

    def GetScaleFactors(self):
        return self._tables[0].copy()

    def GetInverseScaleFactors(self):
        return self._tables[1].copy()

    def GetScaleSigmaSquares(self):
        return self._tables[2].copy()

    def GetInverseScaleSigmaSquares(self):
        return self._tables[3].copy()

    @property
    def mnFeaturesPerLevel(self):
        return self._tables[4].copy()

    def _cap(self):
        return sum(int(q) for q in self._tables[4]) + 64 * self.nlevels

    # ---- operator() (src/ORBextractor.cc:1086-1168) ----
    def __call__(self, image, mask=None, vLappingArea=(0, 1000)):
        """Returns (monoIndex, keypoints[KEYPOINT_DTYPE], descriptors[n,32] uint8); -1 and empty outputs for an empty
        image, like the reference.  `mask` is ignored (include/ORBextractor.h:56)."""
        image = np.asarray(image)
        if image.size == 0:
            return -1, np.empty(0, KEYPOINT_DTYPE), np.empty((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 expected (src/ORBextractor.cc:1094)"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = self._cap()
        kps = np.empty(cap, KEYPOINT_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n, mono = C.c_int(0), C.c_int(0)
        check(self._L.orbx_extract(self._h, _p(image), w, h, image.strides[0], int(vLappingArea[0]), int(vLappingArea[1]),
                                   _p(kps), _p(desc), cap, C.byref(n), C.byref(mono)), self._h, "orbx_extract")
        self._frames = 1
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images, vLappingArea=(0, 1000)):
        """Throughput mode: `images` = array [n,h,w] uint8 (or list of equal-shape 2-D arrays).  Returns a list of
        (monoIndex, keypoints, descriptors) per image."""
        imgs = [np.ascontiguousarray(im, np.uint8) for im in images]
        nimg = len(imgs)
        h, w = imgs[0].shape
        ptrs = (C.c_void_p * nimg)(*[im.ctypes.data for im in imgs])
        cap = self._cap()
        kps = np.empty((nimg, cap), KEYPOINT_DTYPE)
        desc = np.empty((nimg, cap, 32), np.uint8)
        n = np.zeros(nimg, np.int32)
        mono = np.zeros(nimg, np.int32)
        check(self._L.orbx_extract_batch(self._h, ptrs, nimg, w, h, w, int(vLappingArea[0]), int(vLappingArea[1]),
                                         _p(kps), _p(desc), cap, _p(n), _p(mono)), self._h, "orbx_extract_batch")
        self._frames = min(nimg, self.max_batch)
        return [(int(mono[i]), kps[i, :n[i]].copy(), desc[i, :n[i]].copy()) for i in range(nimg)]

    def extract_batch_into(self, images, kps, desc, n_out, mono_out, vLappingArea=(0, 1000)):
        """Lowest-overhead host path: `images` [n,h,w] uint8 C-contiguous, results written into caller-owned arrays
        kps [n,cap] KEYPOINT_DTYPE, desc [n,cap,32] uint8, n_out / mono_out [n] int32.  Page-locked arrays (e.g.
        torch.empty(..., pin_memory=True).numpy()) are used by the library in place, without a staging copy."""
        nimg, h, w = images.shape
        assert images.dtype == np.uint8 and images.flags.c_contiguous
        cap = kps.shape[1]
        base = images.ctypes.data
        ptrs = (C.c_void_p * nimg)(*[base + i * h * w for i in range(nimg)])
        check(self._L.orbx_extract_batch(self._h, ptrs, nimg, w, h, w, int(vLappingArea[0]), int(vLappingArea[1]),
                                         _p(kps), _p(desc), cap, _p(n_out), _p(mono_out)), self._h, "orbx_extract_batch")
        self._frames = min(nimg, self.max_batch)

    # ---- device-resident mode ----
    def extract_device(self, d_ptr, image_bytes, nimg, width, height, stride, vLappingArea=(0, 1000), stream=None):
        check(self._L.orbx_extract_device(self._h, C.c_void_p(d_ptr), image_bytes, nimg, width, height, stride,
                                          int(vLappingArea[0]), int(vLappingArea[1]), C.c_void_p(stream or 0)),
              self._h, "orbx_extract_device")
        self._frames = nimg

    def synchronize(self):
        check(self._L.orbx_synchronize(self._h), self._h, "orbx_synchronize")

    def fetch(self, frame):
        cap = self._cap()
        kps = np.empty(cap, KEYPOINT_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n, mono = C.c_int(0), C.c_int(0)
        check(self._L.orbx_fetch_results(self._h, frame, _p(kps), _p(desc), cap, C.byref(n), C.byref(mono)), self._h,
              "orbx_fetch_results")
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def results_device(self, frame):
        """(device pointer to keypoints, device pointer to descriptors as uint32x8 rows, n, monoIndex)"""
        dk, dd, n, mono = C.c_void_p(), C.c_void_p(), C.c_int(0), C.c_int(0)
        check(self._L.orbx_results_device(self._h, frame, C.byref(dk), C.byref(dd), C.byref(n), C.byref(mono)), self._h,
              "orbx_results_device")
        return dk.value, dd.value, n.value, mono.value

    def results_batch_device(self):
        """(d_kps, d_desc, d_counts, frame_stride): the device arrays holding the last pass's results of all frames."""
        dk, dd, dc, st = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_int(0)
        check(self._L.orbx_results_batch_device(self._h, C.byref(dk), C.byref(dd), C.byref(dc), C.byref(st)), self._h,
              "orbx_results_batch_device")
        return dk.value, dd.value, dc.value, st.value

    # ---- mvImagePyramid (include/ORBextractor.h:83) and stage taps ----
    def level_size(self, level):
        w, h = C.c_int(0), C.c_int(0)
        check(self._L.orbx_level_size(self._h, level, C.byref(w), C.byref(h)), self._h, "orbx_level_size")
        return w.value, h.value

    def image_pyramid_level(self, level, frame=0, padded=False):
        w, h = self.level_size(level)
        if padded:
            w, h = w + 38, h + 38
        out = np.empty((h, w), np.uint8)
        check(self._L.orbx_download_level(self._h, frame, level, int(padded), _p(out), w), self._h, "orbx_download_level")
        return out

    @property
    def mvImagePyramid(self):
        return [self.image_pyramid_level(l) for l in range(self.nlevels)]

    def blurred_level(self, level, frame=0):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        check(self._L.orbx_download_blurred(self._h, frame, level, _p(out), w), self._h, "orbx_download_blurred")
        return out

    def candidates(self, level, frame=0):
        w, h = self.level_size(level)
        cap = (w * h) // 4 + 64
        out = np.empty((cap, 3), np.int32)
        n = C.c_int(0)
        check(self._L.orbx_download_candidates(self._h, frame, level, _p(out), cap, C.byref(n)), self._h,
              "orbx_download_candidates")
        return out[:n.value].copy()

    def level_keypoints(self, level, frame=0):
        cap = self._cap()
        out = np.empty((cap, 3), np.int32)
        n = C.c_int(0)
        check(self._L.orbx_download_level_keypoints(self._h, frame, level, _p(out), cap, C.byref(n)), self._h,
              "orbx_download_level_keypoints")
        return out[:n.value].copy()

    def debug_sort_nodes(self, size, ulx, heap_only=False):
        """heap_only doubles as the mode: 0 / False = one-thread std::sort emulation, 1 / True = heap sort, 2 = warp-parallel form."""
        size = np.ascontiguousarray(size, np.int32)
        ulx = np.ascontiguousarray(ulx, np.int32)
        perm = np.empty(len(size), np.int32)
        check(self._L.orbx_debug_sort_nodes(self._h, _p(size), _p(ulx), len(size), _p(perm), int(heap_only)), self._h,
              "orbx_debug_sort_nodes")
        return perm

    def debug_distribute(self, keys, minX, maxX, minY, maxY, N, pyramid=None):
        """Standalone DistributeOctTree.  pyramid = None: the partitioning form; True: the count-pyramid form first, returns
        (keys, used_pyramid) - used_pyramid is False when a node deeper than the pyramid made the partitioning form take over."""
        keys = np.ascontiguousarray(keys, np.int32).reshape(-1, 3)
        cap = max(N + 64, 4 * 64)
        out = np.empty((cap, 3), np.int32)
        n = C.c_int(0)
        if pyramid is None:
            check(self._L.orbx_debug_distribute(self._h, _p(keys), len(keys), minX, maxX, minY, maxY, N, _p(out), cap,
                                                C.byref(n)), self._h, "orbx_debug_distribute")
            return out[:n.value].copy()
        used = C.c_int(0)
        check(self._L.orbx_debug_distribute2(self._h, _p(keys), len(keys), minX, maxX, minY, maxY, N, int(bool(pyramid)), _p(out), cap,
                                             C.byref(n), C.byref(used)), self._h, "orbx_debug_distribute2")
        return out[:n.value].copy(), bool(used.value)

    # ---- bookkeeping ----
    def launch_count(self):
        return int(self._L.orbx_launch_count(self._h))

    def stream(self):
        return self._L.orbx_stream(self._h)

    def set_stage_timing(self, on=True):
        check(self._L.orbx_set_stage_timing(self._h, int(on)), self._h, "orbx_set_stage_timing")

    def stage_times(self):
        ms = np.zeros(7, np.float32)
        check(self._L.orbx_stage_times(self._h, _p(ms)), self._h, "orbx_stage_times")
        return dict(zip(["pyramid", "fast", "octree", "blur", "finalize", "orient_brief", "total"], ms.tolist()))


def ComputeStereoMatches(left, right, mb, mbf):
    """Frame::ComputeStereoMatches (reference src/Frame.cc:811-981) on the device-resident results of the two extractor
    objects of a stereo Frame (each after one single-frame call).  Returns (kept, mvuRight[N], mvDepth[N])."""
    cap = left._cap()
    ur = np.empty(cap, np.float32)
    dp = np.empty(cap, np.float32)
    n, kept = C.c_int(0), C.c_int(0)
    check(left._L.orbx_stereo_matches(left._h, right._h, float(mb), float(mbf), _p(ur), _p(dp), cap, C.byref(n), C.byref(kept)),
          left._h, "orbx_stereo_matches")
    return kept.value, ur[:n.value].copy(), dp[:n.value].copy()
