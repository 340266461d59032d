#!/usr/bin/env python3
"""The CUDA estimators against cv2 on random two-view scenes, on the GPU box (cv2 is in the image).

tools/ransac_agreement.py replays OpenCV's loop around a HOST build of the minimal solvers (build container, no
GPU); this is the same question asked of the shipped kernels: for scenes over a range of sizes, inlier ratios and
noise, is the inlier mask of
  orbslam2.utils.find_fundamental_ransac      (vslam_ransac_fundamental_host)   == cv2.findFundamentalMat(FM_RANSAC)
  orbslam2.utils.calculate_essential_matrix   (vslam_ransac_essential_host)     == cv2.findEssentialMat(RANSAC)
Prints per estimator the share of identical masks, the worst Jaccard and the largest count difference.
usage: python tools/ransac_agreement_gpu.py [scenes=200]
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))
from orbslam2 import utils as U  # noqa: E402


def scene(rng, n, inlier_ratio, noise, K):
    R, _ = cv2.Rodrigues(rng.normal(0, 0.08, 3))
    t = rng.normal(0, 0.3, 3)
    X = rng.uniform(-2, 2, (n, 3)) + np.array([0, 0, 6.0])
    x1 = (K @ X.T).T
    x2 = (K @ ((R @ X.T).T + t).T).T
    p1 = x1[:, :2] / x1[:, 2:] + rng.normal(0, noise, (n, 2))
    p2 = x2[:, :2] / x2[:, 2:] + rng.normal(0, noise, (n, 2))
    out = rng.random(n) > inlier_ratio
    p2[out] = rng.uniform(0, 640, (int(out.sum()), 2))
    return p1.astype(np.float32), p2.astype(np.float32)


def jacc(a, b):
    u = (a | b).sum()
    return 1.0 if u == 0 else (a & b).sum() / u


def main():
    n_scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    K = np.array([[500.0, 0, 320], [0, 500, 240], [0, 0, 1]])
    rng = np.random.default_rng(11)
    stats = {}
    for s in range(n_scenes):
        n = int(rng.integers(40, 2000))
        ratio = float(rng.uniform(0.3, 0.95))
        noise = float(rng.uniform(0.1, 1.0))
        p1, p2 = scene(rng, n, ratio, noise, K)
        _, mc = cv2.findFundamentalMat(p1, p2, cv2.FM_RANSAC, 3.0, 0.99)
        _, mp = U.find_fundamental_ransac(p1, p2, 3.0, 0.99)
        mc = np.zeros(n, bool) if mc is None else mc.ravel().astype(bool)
        mp = mp.ravel().astype(bool)
        stats.setdefault("F (thr 3, conf 0.99)", []).append((np.array_equal(mp, mc), jacc(mp, mc), abs(int(mp.sum()) - int(mc.sum()))))
        for thr in (1.0, 3.0):
            _, mc = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, thr)
            _, mp = U.calculate_essential_matrix(p1, p2, K, cv2.RANSAC, 0.999, thr)
            mc = np.zeros(n, bool) if mc is None else mc.ravel().astype(bool)
            mp = mp.ravel().astype(bool)
            stats.setdefault("E (thr %g, prob 0.999)" % thr, []).append((np.array_equal(mp, mc), jacc(mp, mc), abs(int(mp.sum()) - int(mc.sum()))))
    print("# CUDA estimators vs cv2 %s on %d random scenes (40-2000 points, 30-95 %% inliers, noise 0.1-1 px)" % (cv2.__version__, n_scenes))
    for k, v in stats.items():
        eq = np.mean([x[0] for x in v])
        js = np.array([x[1] for x in v])
        print("%-24s identical masks %5.1f %%   Jaccard min %.4f  mean %.6f   max |count difference| %d" %
              (k, 100 * eq, js.min(), js.mean(), max(x[2] for x in v)))


if __name__ == "__main__":
    main()
