#!/usr/bin/env python3
"""Generates tests/golden/cv2_undistort.npz with the REAL cv2.undistortPoints (cv2 4.13.0 in the build container): the
known-answer vectors for Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:747-812), which call
cv::undistortPoints(mat, mat, K, distCoef, Mat(), K).  Run from the repo root:  python tools/gen_golden_undistort.py
"""
import os
import numpy as np
import cv2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rng = np.random.Generator(np.random.PCG64(20261105))
CAMS = {
    # EuRoC cam0 (Examples/Monocular/EuRoC.yaml of the upstream project), 4 coefficients
    "euroc": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05), (752, 480)),
    # TUM1: 5 coefficients with a large k3
    "tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314), (640, 480)),
    # KITTI-like: zero tangential terms, tiny k1
    "smallk": ((718.856, 718.856, 607.1928, 185.2157), (1e-4, 0.0, 0.0, 0.0), (1241, 376)),
    # rational model, 8 coefficients
    "rational": ((600.0, 610.0, 320.5, 240.25), (0.12, -0.05, 0.001, -0.0007, 0.01, 0.08, -0.02, 0.003), (640, 480)),
    # thin prism, 12 coefficients
    "prism": ((420.0, 421.5, 330.0, 250.0), (-0.2, 0.04, 0.0005, -0.0003, 0.002, 0.01, 0.0, 0.0, 0.001, -0.0004, 0.0007, 0.0002), (640, 480)),
    # strong barrel distortion far outside the calibrated range: exercises the icdist < 0 escape
    "wild": ((200.0, 200.0, 320.0, 240.0), (-0.9, 0.1, 0.0, 0.0, -0.5), (640, 480)),
}
out = {}
for name, (K4, dist, (w, h)) in CAMS.items():
    K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
    d = np.array(dist, np.float32)
    xy = np.stack([rng.uniform(0, w, 1500), rng.uniform(0, h, 1500)], 1).astype(np.float32)
    xy[:700] = np.round(xy[:700])                      # level-0 keypoints are integers; upper levels are not
    corners = np.array([[0, 0], [w, 0], [0, h], [w, h]], np.float32)     # ComputeImageBounds
    xy = np.concatenate([corners, xy])
    ref = cv2.undistortPoints(xy.reshape(-1, 1, 2).copy(), K, d, None, K).reshape(-1, 2)
    out[name + "_K"] = np.array(K4, np.float32)
    out[name + "_dist"] = d
    out[name + "_xy"] = xy
    out[name + "_out"] = ref.astype(np.float32)
np.savez_compressed(os.path.join(ROOT, "tests/golden/cv2_undistort.npz"), **out)
print("wrote", len(out), "arrays")
