"""ORACLE — TEST INFRASTRUCTURE ONLY.

The reference's pixel stages driven with the REAL OpenCV (cv2 4.13, the only OpenCV in this image): pyramid
(cv2.resize INTER_LINEAR + copyMakeBorder REFLECT_101), per-cell cv2 FAST with the iniTh/minTh fallback,
cv2.GaussianBlur 7x7 sigma 2 and cv2.fastAtan2.  Control flow follows src/ORBextractor.cc:781-872, 1170-1195.
Used only to pin oracle/cpp (tests/test_oracle_vs_cv2.py) and to generate tests/golden/ (tools/gen_golden.py);
cv2 does not travel to the parity tests on the GPU box as a dependency of the product.
"""
import math
import numpy as np
import cv2

EDGE = 19


def scale_tables(scale_factor, nlevels):
    """float32 tables exactly as the ctor builds them (src/ORBextractor.cc:414-431)."""
    sf = float(np.float32(scale_factor))           # float argument stored in a double member
    sc = [np.float32(1.0)]
    for _ in range(1, nlevels):
        sc.append(np.float32(float(sc[-1]) * sf))
    inv = [np.float32(1.0) / s for s in sc]
    return np.array(sc, np.float32), np.array(inv, np.float32)


def cv_round(x):
    return int(np.rint(x))


def pyramid(img, scale_factor=1.2, nlevels=8):
    """Returns list of (padded, roi_view) per level (src/ORBextractor.cc:1170-1195)."""
    _, inv = scale_tables(scale_factor, nlevels)
    h, w = img.shape
    out = []
    for l in range(nlevels):
        s = inv[l]
        sw = cv_round(np.float32(w) * s)
        sh = cv_round(np.float32(h) * s)
        if l == 0:
            padded = cv2.copyMakeBorder(img, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
        else:
            prev = out[-1][1]
            r = cv2.resize(prev, (sw, sh), interpolation=cv2.INTER_LINEAR)
            padded = cv2.copyMakeBorder(r, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
        out.append((padded, padded[EDGE:EDGE + sh, EDGE:EDGE + sw]))
    return out


def fast_cells(level_img, ini_th=20, min_th=7):
    """Per-cell FAST candidates of one level, coords relative to minBorder (src/ORBextractor.cc:787-868).
    Returns (array n x 3 of x, y, response; number of cv2.FAST calls; number of minTh fallbacks)."""
    h, w = level_img.shape
    minB = EDGE - 3
    maxBX, maxBY = w - EDGE + 3, h - EDGE + 3
    width, height = np.float32(maxBX - minB), np.float32(maxBY - minB)
    W = np.float32(35)
    nCols, nRows = int(width / W), int(height / W)
    wCell, hCell = int(math.ceil(width / np.float32(nCols))), int(math.ceil(height / np.float32(nRows)))
    det_ini = cv2.FastFeatureDetector_create(ini_th, True)
    det_min = cv2.FastFeatureDetector_create(min_th, True)
    pts = []
    calls = fallbacks = 0
    for i in range(nRows):
        iniY = minB + i * hCell
        maxY = iniY + hCell + 6
        if iniY >= maxBY - 3:
            continue
        maxY = min(maxY, maxBY)
        for j in range(nCols):
            iniX = minB + j * wCell
            maxX = iniX + wCell + 6
            if iniX >= maxBX - 6:
                continue
            maxX = min(maxX, maxBX)
            cell = np.ascontiguousarray(level_img[iniY:maxY, iniX:maxX])
            kps = det_ini.detect(cell)
            calls += 1
            if len(kps) == 0:
                kps = det_min.detect(cell)
                calls += 1
                fallbacks += 1
            for k in kps:
                pts.append((k.pt[0] + j * wCell, k.pt[1] + i * hCell, k.response))
    return np.array(pts, np.float32).reshape(-1, 3), calls, fallbacks


def blur(level_img):
    return cv2.GaussianBlur(np.ascontiguousarray(level_img), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)


def synth_frame(w=752, h=480, seed=0, shift=(0.0, 0.0), rot_deg=0.0):
    """Synthetic textured grayscale frame: multi-octave box-blurred noise + random
    rectangles/discs for strong corners + a flat band so some cells hit the minThFAST fallback."""
    rng = np.random.Generator(np.random.PCG64(seed))
    big = 64
    H, W = h + 2 * big, w + 2 * big
    acc = np.zeros((H, W), np.float32)
    for k in (1, 2, 4, 8, 16):
        n = rng.standard_normal((H, W)).astype(np.float32)
        if k > 1:
            n = cv2.blur(n, (k, k))
            n *= k
        acc += n
    acc -= acc.min()
    acc *= 255.0 / acc.max()
    img = acc.astype(np.uint8)
    for _ in range(300):
        x, y = int(rng.integers(0, W)), int(rng.integers(0, H))
        c = int(rng.integers(0, 256))
        if rng.random() < 0.5:
            cv2.rectangle(img, (x, y), (x + int(rng.integers(6, 60)), y + int(rng.integers(6, 60))), c, -1)
        else:
            cv2.circle(img, (x, y), int(rng.integers(4, 30)), c, -1)
    # flat / low-contrast band: ~1/8 of the width
    bx = big + w // 2
    band = img[:, bx:bx + w // 8].astype(np.float32)
    img[:, bx:bx + w // 8] = (128 + (band - 128) * 0.04).astype(np.uint8)
    img[big + h // 3: big + h // 3 + h // 6, big: big + w // 6] = 90
    if shift != (0.0, 0.0) or rot_deg != 0.0:
        M = cv2.getRotationMatrix2D((W / 2, H / 2), rot_deg, 1.0)
        M[0, 2] += shift[0]
        M[1, 2] += shift[1]
        img = cv2.warpAffine(img, M, (W, H), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT_101)
    return np.ascontiguousarray(img[big:big + h, big:big + w])
