#!/usr/bin/env python3
"""Generates tests/golden/*.npz with the REAL OpenCV (cv2 4.13.0 in the build container) — the known-answer vectors
that pin oracle/cpp (and, through it, the CUDA path).  The reference ships no golden vectors for this path
(SURVEY.md 4, 8c); its pixel arithmetic is OpenCV's, so cv2 outputs are the closest thing to reference outputs that
can be produced here.  Run from the repo root:  python tools/gen_golden.py
"""
import os
import sys
import numpy as np
import cv2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.py import cv2_pipeline as cvp   # noqa: E402
from mam3slam_b200 import synth            # noqa: E402

cv2.setNumThreads(1)
rng = np.random.Generator(np.random.PCG64(20261104))
out = {}

# ---- resize INTER_LINEAR (src/ORBextractor.cc:1183) ----
cases = [(75, 48, 63, 40), (94, 61, 78, 51), (64, 64, 32, 32), (120, 33, 100, 28), (37, 91, 31, 76)]
for i, (sw, sh, dw, dh) in enumerate(cases):
    src = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
    out[f"resize{i}_src"] = src
    out[f"resize{i}_dst"] = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
out["resize_n"] = np.int32(len(cases))

# ---- copyMakeBorder REFLECT_101 (:1185,1190) ----
src = rng.integers(0, 256, (23, 31), dtype=np.uint8)
out["border_src"] = src
out["border_dst"] = cv2.copyMakeBorder(src, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)

# ---- GaussianBlur 7x7 sigma 2 (:1133) ----
src = rng.integers(0, 256, (64, 80), dtype=np.uint8)
out["blur_src"] = src
out["blur_dst"] = cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
tex = synth.frame(200, 150, 5)
out["blur2_src"] = tex
out["blur2_dst"] = cv2.GaussianBlur(tex, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)

# ---- FAST-9/16 + NMS (:826,845) ----
cells = []
for i in range(8):
    if i < 4:
        c = rng.integers(0, 256, (41, 41), dtype=np.uint8)
    else:
        x, y = int(rng.integers(0, 150)), int(rng.integers(0, 100))
        c = np.ascontiguousarray(tex[y:y + 38 + i, x:x + 44 - i])
    cells.append(c)
    out[f"fast{i}_img"] = c
    for th in (20, 7):
        kps = cv2.FastFeatureDetector_create(th, True).detect(c)
        out[f"fast{i}_th{th}"] = np.array([(k.pt[0], k.pt[1], k.response) for k in kps], np.int32).reshape(-1, 3)
out["fast_n"] = np.int32(len(cells))

# ---- fastAtan2 (:102) ----
yx = np.concatenate([rng.integers(-200000, 200000, (1500, 2)).astype(np.float32),
                     rng.standard_normal((400, 2)).astype(np.float32),
                     np.array([[0, 1], [1, 0], [0, -1], [-1, 0], [1, 1], [-1, -1], [0, 0], [5, 5], [-3, 3]], np.float32)])
out["atan_yx"] = yx
out["atan_deg"] = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in yx], np.float32)

# ---- IC_Angle + rBRIEF through cv2.ORB (same bit_pattern_31_, same ICAngles, same 7x7 blur) ----
img = synth.frame(240, 180, 9)
orb = cv2.ORB_create(nfeatures=400, scaleFactor=1.2, nlevels=1, edgeThreshold=19, firstLevel=0, WTA_K=2,
                     scoreType=cv2.ORB_FAST_SCORE, patchSize=31, fastThreshold=20)
kps = orb.detect(img)
kps2, desc = orb.compute(img, kps)
out["orb_img"] = img
out["orb_kps"] = np.array([(k.pt[0], k.pt[1], k.angle) for k in kps2], np.float32).reshape(-1, 3)
out["orb_desc"] = desc
# cv2.ORB blurs a SUBMATRIX of its pyramid buffer, which makes cv::GaussianBlur skip its bit-exact fixed-point path and
# use the float kernel; the reference blurs a clone() (src/ORBextractor.cc:1132-1133) = the fixed-point path that
# blur_dst above pins.  The float-blurred plane is stored so that the rBRIEF sampling arithmetic can be pinned alone.
k = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
out["orb_blur_float"] = cv2.sepFilter2D(img, cv2.CV_8U, k, k, borderType=cv2.BORDER_REFLECT_101)
np.savez_compressed(os.path.join(ROOT, "tests/golden/cv2_primitives.npz"), **out)

# ---- the reference's pixel stages driven with cv2 on one small frame (control flow of :781-872, :1170-1195) ----
pipe = {}
frame = synth.frame(320, 240, 3)
pipe["frame"] = frame
pyr = cvp.pyramid(frame, 1.2, 4)
for l, (padded, roi) in enumerate(pyr):
    pipe[f"padded{l}"] = padded
    cand, calls, fb = cvp.fast_cells(np.ascontiguousarray(roi), 20, 7)
    pipe[f"cand{l}"] = cand.astype(np.int32)
    pipe[f"blur{l}"] = cvp.blur(roi)
    pipe[f"calls{l}"] = np.int32([calls, fb])
np.savez_compressed(os.path.join(ROOT, "tests/golden/cv2_pipeline_320x240.npz"), **pipe)
for f in ("cv2_primitives.npz", "cv2_pipeline_320x240.npz"):
    print(f, os.path.getsize(os.path.join(ROOT, "tests/golden", f)))
print("orb keypoints:", len(kps2), "cand per level:", [len(pipe[f"cand{l}"]) for l in range(4)])
