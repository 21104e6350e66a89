"""Synthetic textured grayscale frames (numpy only) for tests and bench.py — SURVEY.md 8(d) config 2/3.

Base texture = seeded PCG64 multi-octave noise (box-blurred white noise at 1,2,4,8,16 px, contrast-stretched) +
random filled rectangles / discs for strong corners + a low-contrast band and a flat patch so that some FAST cells
fall back to minThFAST and some stay empty.  A stream is the base warped by a smooth random-walk translation
(<= 8 px) and rotation (<= 2 deg).  There is no dataset on the box: `data` = "synthetic".
"""
import numpy as np


def _box(a, k):
    if k <= 1:
        return a
    pad = k // 2
    p = np.pad(a, ((pad, k - 1 - pad), (pad, k - 1 - pad)), mode="reflect")
    c = np.cumsum(np.cumsum(p, axis=0, dtype=np.float64), axis=1)
    c = np.pad(c, ((1, 0), (1, 0)))
    s = c[k:, k:] - c[:-k, k:] - c[k:, :-k] + c[:-k, :-k]
    return (s / (k * k)).astype(np.float32)


def base_texture(w, h, seed, margin=64, octaves=(1, 2, 4, 8, 16)):
    """octaves: box sizes of the noise components.  The default (with the 1- and 2-px components) is a DENSE texture: ~26 k
    FAST candidates per 752x480 frame; octaves=(8, 16, 32) leaves smooth shading + the shapes' edges, a SPARSE texture closer
    to indoor imagery (a few thousand candidates, most cells decided by a handful of strong corners)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    H, W = h + 2 * margin, w + 2 * margin
    acc = np.zeros((H, W), np.float32)
    for k in octaves:
        acc += _box(rng.standard_normal((H, W)).astype(np.float32), k) * k
    acc -= acc.min()
    acc *= 255.0 / acc.max()
    img = acc.astype(np.uint8)
    yy, xx = np.mgrid[0:H, 0:W]
    for _ in range(300):
        x, y = int(rng.integers(0, W)), int(rng.integers(0, H))
        c = int(rng.integers(0, 256))
        if rng.random() < 0.5:
            img[y:y + int(rng.integers(6, 60)), x:x + int(rng.integers(6, 60))] = c
        else:
            r = int(rng.integers(4, 30))
            y0, y1, x0, x1 = max(y - r, 0), min(y + r + 1, H), max(x - r, 0), min(x + r + 1, W)
            m = (yy[y0:y1, x0:x1] - y) ** 2 + (xx[y0:y1, x0:x1] - x) ** 2 <= r * r
            img[y0:y1, x0:x1][m] = c
    bx = margin + w // 2
    band = img[:, bx:bx + w // 8].astype(np.float32)
    img[:, bx:bx + w // 8] = (128 + (band - 128) * 0.04).astype(np.uint8)
    img[margin + h // 3: margin + h // 3 + h // 6, margin: margin + w // 6] = 90
    return img


def warp(base, w, h, dx, dy, rot_deg, margin=64):
    """Bilinear resample of `base` (with margin) under a rotation about the centre + translation; crop to w x h."""
    H, W = base.shape
    a = np.deg2rad(rot_deg)
    ca, sa = np.cos(a), np.sin(a)
    ys, xs = np.mgrid[margin:margin + h, margin:margin + w].astype(np.float32)
    cx, cy = W / 2.0, H / 2.0
    sx = ca * (xs - cx) - sa * (ys - cy) + cx + dx
    sy = sa * (xs - cx) + ca * (ys - cy) + cy + dy
    sx = np.clip(sx, 0, W - 1.001)
    sy = np.clip(sy, 0, H - 1.001)
    x0, y0 = np.floor(sx).astype(np.int32), np.floor(sy).astype(np.int32)
    fx, fy = sx - x0, sy - y0
    b = base.astype(np.float32)
    v = (b[y0, x0] * (1 - fx) * (1 - fy) + b[y0, x0 + 1] * fx * (1 - fy) +
         b[y0 + 1, x0] * (1 - fx) * fy + b[y0 + 1, x0 + 1] * fx * fy)
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


SPARSE_OCTAVES = (8, 16, 32)


def frame(w=752, h=480, seed=0, sparse=False):
    octs = SPARSE_OCTAVES if sparse else (1, 2, 4, 8, 16)
    return np.ascontiguousarray(base_texture(w, h, seed, octaves=octs)[64:64 + h, 64:64 + w])


def stream(w=752, h=480, seed=0, nframes=8, sparse=False):
    """[nframes, h, w] uint8 frames of one agent: random-walk motion over one base texture."""
    base = base_texture(w, h, seed, octaves=SPARSE_OCTAVES if sparse else (1, 2, 4, 8, 16))
    rng = np.random.Generator(np.random.PCG64(seed + 100003))
    dx = dy = rot = 0.0
    out = np.empty((nframes, h, w), np.uint8)
    for t in range(nframes):
        if t:
            dx = float(np.clip(dx + rng.normal(0, 2.0), -8, 8))
            dy = float(np.clip(dy + rng.normal(0, 2.0), -8, 8))
            rot = float(np.clip(rot + rng.normal(0, 0.5), -2, 2))
        out[t] = warp(base, w, h, dx, dy, rot) if t else base[64:64 + h, 64:64 + w]
    return out


def stereo_pair(w=752, h=480, seed=0, disparity=17.3, noise=2):
    """(left, right) of a rectified pair: the right image sees the base texture shifted by a sub-pixel disparity (a
    fronto-parallel plane) plus a little sensor noise, so that Frame::ComputeStereoMatches has matches to refine."""
    base = base_texture(w, h, seed)
    left = np.ascontiguousarray(base[64:64 + h, 64:64 + w])
    right = warp(base, w, h, float(disparity), 0.0, 0.0).astype(np.int16)
    rng = np.random.Generator(np.random.PCG64(seed + 7919))
    right += rng.integers(-noise, noise + 1, right.shape, dtype=np.int16)
    return left, np.clip(right, 0, 255).astype(np.uint8)
