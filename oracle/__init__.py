"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes bindings of the CPU restatement (oracle/cpp).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package; the product (mam3slam_b200/) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    """Compile liborb_oracle.so with the committed Makefile (gcc only)."""
    so = os.path.join(_HERE, "liborb_oracle.so")
    if force and os.path.exists(so):
        os.remove(so)
    subprocess.check_call(["make", "-s", "-C", _HERE, "liborb_oracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liborb_oracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        u8p, i32p, f32p = C.POINTER(C.c_uint8), C.POINTER(C.c_int), C.POINTER(C.c_float)
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_extractor_destroy.argtypes = [C.c_void_p]
        L.orc_extract.restype = C.c_int
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_int, i32p]
        L.orc_extract_timed.restype = C.c_double
        L.orc_extract_timed.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.POINTER(C.c_long)]
        L.orc_extractor_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orc_level_size.argtypes = [C.c_void_p, C.c_int, i32p, i32p]
        L.orc_level_pixels.restype = C.c_int
        L.orc_level_pixels.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_level_points.restype = C.c_int
        L.orc_level_points.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_sort_tie_count.restype = C.c_long
        L.orc_sort_tie_count.argtypes = [C.c_void_p]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_fast9_nms.restype = C.c_int
        L.orc_fast9_nms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_resize_linear.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_border101.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.orc_gaussian7.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_sort_nodes.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_distribute.restype = C.c_int
        L.orc_distribute.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_void_p, C.c_int]
        L.orc_ic_angles.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_brief.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# ---- primitives ---------------------------------------------------------------------------------------------
def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), dw, dh, dw)
    return dst


def border101(src, b):
    src = np.ascontiguousarray(src, np.uint8)
    h, w = src.shape
    dst = np.empty((h + 2 * b, w + 2 * b), np.uint8)
    lib().orc_border101(_p(src), w, h, w, _p(dst), w + 2 * b, b)
    return dst


def gaussian7(src):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty_like(src)
    lib().orc_gaussian7(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), src.shape[1])
    return dst


def fast_atan2(y, x):
    return float(lib().orc_fast_atan2(float(y), float(x)))


def fast9_nms(img, th):
    img = np.ascontiguousarray(img, np.uint8)
    cap = img.size // 4 + 16
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_fast9_nms(_p(img), img.shape[1], img.shape[0], img.shape[1], int(th), _p(out), cap)
    return out[:n].copy()


def sort_nodes(size, ulx):
    size = np.ascontiguousarray(size, np.int32)
    ulx = np.ascontiguousarray(ulx, np.int32)
    perm = np.empty(len(size), np.int32)
    lib().orc_sort_nodes(_p(size), _p(ulx), len(size), _p(perm))
    return perm


def distribute(keys, minX, maxX, minY, maxY, N):
    keys = np.ascontiguousarray(keys, np.int32).reshape(-1, 3)
    out = np.empty((max(4 * N + 64, 64), 3), np.int32)
    n = lib().orc_distribute(0, _p(keys), len(keys), minX, maxX, minY, maxY, N, _p(out), len(out))
    return out[:n].copy()


def ic_angles(img, xy):
    """IC_Angle (degrees) at integer positions xy[n,2] of a plain image."""
    img = np.ascontiguousarray(img, np.uint8)
    xy = np.ascontiguousarray(xy, np.int32).reshape(-1, 2)
    out = np.empty(len(xy), np.float32)
    lib().orc_ic_angles(_p(img), img.shape[1], img.shape[0], img.shape[1], _p(xy), len(xy), _p(out))
    return out


def brief(blurred, kps):
    """256-bit rotated-BRIEF rows for kps[n,3] = (x, y, angle_deg) on an already blurred plain image."""
    blurred = np.ascontiguousarray(blurred, np.uint8)
    kps = np.ascontiguousarray(kps, np.float32).reshape(-1, 3)
    out = np.empty((len(kps), 32), np.uint8)
    lib().orc_brief(_p(blurred), blurred.shape[1], blurred.shape[0], blurred.shape[1], _p(kps), len(kps), _p(out))
    return out


# ---- extractor ------------------------------------------------------------------------------------------------
class Extractor:
    """CPU oracle with the reference's ORBextractor constructor arguments (src/ORBextractor.cc:409)."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7):
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self._h = lib().orc_extractor_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h)
            self._h = None

    def tables(self):
        n = self.nlevels
        f = [np.empty(n, np.float32) for _ in range(4)]
        per = np.empty(n, np.int32)
        umax = np.empty(16, np.int32)
        lib().orc_extractor_tables(self._h, _p(f[0]), _p(f[1]), _p(f[2]), _p(f[3]), _p(per), _p(umax))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], per_level=per, umax=umax)

    def __call__(self, image, lapping=(0, 1000)):
        """Returns (monoIndex, keypoints[n,6] float32 = x,y,size,angle,response,octave, descriptors[n,32] u8)."""
        image = np.ascontiguousarray(image, np.uint8)
        h, w = image.shape
        cap = self.nfeatures * 2 + 64 * self.nlevels
        kps = np.empty((cap, 6), np.float32)
        desc = np.empty((cap, 32), np.uint8)
        n = C.c_int(0)
        mono = lib().orc_extract(self._h, _p(image), w, h, w, int(lapping[0]), int(lapping[1]), _p(kps), _p(desc), cap,
                                 C.byref(n))
        assert n.value <= cap
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def timed(self, images, iters, lapping=(0, 1000)):
        images = np.ascontiguousarray(images, np.uint8)
        nimg, h, w = images.shape
        tot = C.c_long(0)
        sec = lib().orc_extract_timed(self._h, _p(images), nimg, w, h, w, int(lapping[0]), int(lapping[1]), iters,
                                      C.byref(tot))
        return sec, tot.value

    def level_size(self, level):
        w, h = C.c_int(0), C.c_int(0)
        lib().orc_level_size(self._h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_pixels(self, level, which="plain"):
        w, h = self.level_size(level)
        if which == "padded":
            out = np.empty((h + 38, w + 38), np.uint8)
            lib().orc_level_pixels(self._h, level, 1, _p(out))
        elif which == "blurred":
            out = np.empty((h, w), np.uint8)
            if lib().orc_level_pixels(self._h, level, 2, _p(out)) == 0:
                return None
        else:
            out = np.empty((h, w), np.uint8)
            lib().orc_level_pixels(self._h, level, 0, _p(out))
        return out

    def level_points(self, level, kept=False):
        cap = 1 << 18
        out = np.empty((cap, 4), np.float32)
        n = lib().orc_level_points(self._h, level, 1 if kept else 0, _p(out), cap)
        assert n <= cap
        return out[:n].copy()

    def sort_tie_count(self):
        return int(lib().orc_sort_tie_count(self._h))
