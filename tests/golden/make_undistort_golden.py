"""Generates tests/golden/undistort.npz: outputs of cv2.undistort (what utils.undistort_image of the reference
returns, /root/reference/src/orbslam2/utils.py:40-52) on small seeded images.  Run in the build container:
    python tests/golden/make_undistort_golden.py
"""
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, "/root/reference/src")
from orbslam2.utils import undistort_image   # the reference's own wrapper

rng = np.random.default_rng(2026)
out = {"cv2_version": np.array(cv2.__version__)}
cases = [(97, 131, 1, 5), (120, 160, 3, 5), (64, 4200, 1, 4), (150, 90, 1, 8), (80, 100, 1, 12), (60, 70, 1, 0)]
for n, (h, w, ch, nd) in enumerate(cases):
    shape = (h, w) if ch == 1 else (h, w, ch)
    img = cv2.GaussianBlur(rng.integers(0, 256, shape, dtype=np.uint8), (0, 0), 1.2)
    K = np.array([[0.8 * w + rng.uniform(-9, 9), 0, w / 2 + rng.uniform(-7, 7)],
                  [0, 0.8 * w + rng.uniform(-9, 9), h / 2 + rng.uniform(-7, 7)], [0, 0, 1]])
    full = np.array([rng.uniform(-0.3, 0.3), rng.uniform(-0.1, 0.1), rng.uniform(-0.005, 0.005),
                     rng.uniform(-0.005, 0.005), rng.uniform(-0.05, 0.05), rng.uniform(-0.05, 0.05),
                     rng.uniform(-0.02, 0.02), rng.uniform(-0.01, 0.01), rng.uniform(-0.003, 0.003),
                     rng.uniform(-0.003, 0.003), rng.uniform(-0.003, 0.003), rng.uniform(-0.003, 0.003)])
    dist = full[:nd] if nd else np.zeros(5)
    out["img%d" % n], out["K%d" % n], out["dist%d" % n] = img, K, dist
    out["out%d" % n] = undistort_image(img, K, dist)
out["n"] = np.array(len(cases))
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "undistort.npz"), **out)
print("wrote undistort.npz", len(cases), "cases")
