"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes bindings of oracle/_ref/liborb_ref*.so: the reference's OWN src/ORBextractor.cc compiled unmodified against the
header shim oracle/ref_shim/ (recipe: oracle/Makefile, target `ref`; entry points: oracle/ref_capi.cc).  It needs
/root/reference to BUILD; the built libraries are git-ignored and travel to the GPU box with the snapshot, so tests use
`available()` and skip when neither the library nor the reference checkout is there.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {}
REF_ROOT = os.environ.get("MAM3SLAM_REFERENCE", "/root/reference")


def _so(fma):
    return os.path.join(_HERE, "_ref", "liborb_ref_fma.so" if fma else "liborb_ref.so")


def build():
    """Compile the reference's extractor where it lies (only possible where the reference checkout exists)."""
    if not os.path.exists(os.path.join(REF_ROOT, "src", "ORBextractor.cc")):
        return False
    subprocess.check_call(["make", "-s", "-C", _HERE, "ref", f"REF={REF_ROOT}"])
    return True


def available(fma=False):
    return os.path.exists(_so(fma)) or build()


def lib(fma=False):
    if fma not in _LIBS:
        if not available(fma):
            raise RuntimeError("oracle/_ref is not built and the reference checkout is absent")
        L = C.CDLL(_so(fma))
        i32p = C.POINTER(C.c_int)
        L.ref_extractor_create.restype = C.c_void_p
        L.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_extractor_destroy.argtypes = [C.c_void_p]
        L.ref_extractor_tables.argtypes = [C.c_void_p] * 7
        L.ref_extract.restype = C.c_int
        L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_int, i32p]
        L.ref_extract_timed.restype = C.c_double
        L.ref_extract_timed.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.POINTER(C.c_long)]
        L.ref_level_size.argtypes = [C.c_void_p, C.c_int, i32p, i32p]
        L.ref_level_pixels.restype = C.c_int
        L.ref_level_pixels.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ref_level_keypoints.restype = C.c_int
        L.ref_level_keypoints.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.ref_distribute.restype = C.c_int
        L.ref_distribute.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        _LIBS[fma] = L
    return _LIBS[fma]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def distribute(keys, minX, maxX, minY, maxY, N):
    """ORBextractor::DistributeOctTree (src/ORBextractor.cc:555-779) on keys[n,3] = (x, y, response)."""
    keys = np.ascontiguousarray(keys, np.int32).reshape(-1, 3)
    out = np.empty((max(4 * N + 64, 64), 3), np.int32)
    n = lib().ref_distribute(_p(keys), len(keys), minX, maxX, minY, maxY, N, _p(out), len(out))
    return out[:n].copy()


class Extractor:
    """MAM3SLAM::ORBextractor itself (src/ORBextractor.cc:409), same Python surface as oracle.Extractor."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, fma=False):
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self._L = lib(fma)
        self._h = self._L.ref_extractor_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.ref_extractor_destroy(self._h)
            self._h = None

    def tables(self):
        n = self.nlevels
        f = [np.empty(n, np.float32) for _ in range(4)]
        per = np.empty(n, np.int32)
        umax = np.empty(16, np.int32)
        self._L.ref_extractor_tables(self._h, _p(f[0]), _p(f[1]), _p(f[2]), _p(f[3]), _p(per), _p(umax))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], per_level=per, umax=umax)

    def __call__(self, image, lapping=(0, 1000)):
        """Returns (monoIndex, keypoints[n,6] float32 = x,y,size,angle,response,octave, descriptors[n,32] u8)."""
        image = np.ascontiguousarray(image, np.uint8)
        h, w = image.shape
        cap = self.nfeatures * 2 + 64 * self.nlevels
        kps = np.empty((cap, 6), np.float32)
        desc = np.empty((cap, 32), np.uint8)
        n = C.c_int(0)
        mono = self._L.ref_extract(self._h, _p(image), w, h, w, int(lapping[0]), int(lapping[1]), _p(kps), _p(desc), cap,
                                   C.byref(n))
        assert n.value <= cap
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def timed(self, images, iters, lapping=(0, 1000)):
        images = np.ascontiguousarray(images, np.uint8)
        nimg, h, w = images.shape
        tot = C.c_long(0)
        sec = self._L.ref_extract_timed(self._h, _p(images), nimg, w, h, w, int(lapping[0]), int(lapping[1]), iters,
                                        C.byref(tot))
        return sec, tot.value

    def level_size(self, level):
        w, h = C.c_int(0), C.c_int(0)
        self._L.ref_level_size(self._h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_pixels(self, level, which="plain"):
        w, h = self.level_size(level)
        b = 19 if which == "padded" else 0
        out = np.empty((h + 2 * b, w + 2 * b), np.uint8)
        self._L.ref_level_pixels(self._h, level, 1 if b else 0, _p(out))
        return out

    def level_keypoints(self, image, level):
        """Kept keypoints of one level after ComputeKeyPointsOctTree: [n,4] = x, y, response, angle (level coords)."""
        cap = 1 << 16
        out = np.empty((cap, 4), np.float32)
        if image is not None:
            image = np.ascontiguousarray(image, np.uint8)
            h, w = image.shape
            n = self._L.ref_level_keypoints(self._h, _p(image), w, h, w, level, _p(out), cap)
        else:
            n = self._L.ref_level_keypoints(self._h, None, 0, 0, 0, level, _p(out), cap)
        return out[:n].copy()
