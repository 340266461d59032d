#!/usr/bin/env python3
"""Golden vectors for the RANSAC estimators, made by RUNNING THE REFERENCE (build container only).

Synthetic two-view scenes (seeded; 40..1500 correspondences, 30..95 % inliers, 0.1..1 px noise, so that the
RANSAC loops run for tens to hundreds of iterations -- the three image-pair goldens stop after 2..9) through
  * DescriptorMatcher.filter_matches_by_fundamental (src/orbslam2/matcher.py:171-200) at threshold 3,
  * utils.calculate_essential_matrix (src/orbslam2/utils.py:120-126) at thresholds 1 and 3 (prob 0.999),
and small sets (8..14 matches) where cv2 switches to LMedS.  Stored: the float32 points, K, and the masks /
matrices the reference returned.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_ransac_golden.py
"""
import os
import sys

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "src"))

import cv2  # noqa: E402
import numpy as np  # noqa: E402

from orbslam2.matcher import DescriptorMatcher  # noqa: E402  (the reference's)
from orbslam2 import utils as ref_utils  # noqa: E402

K = np.array([[500.0, 0, 320], [0, 500, 240], [0, 0, 1]])
N_SCENES = 24


def scene(rng, n, inlier_ratio, noise):
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


def fundamental_through_reference(p1, p2):
    kp1 = [cv2.KeyPoint(float(x), float(y), 31) for x, y in p1]
    kp2 = [cv2.KeyPoint(float(x), float(y), 31) for x, y in p2]
    matches = [cv2.DMatch(i, i, 0, 0.0) for i in range(len(p1))]
    _, mask = DescriptorMatcher().filter_matches_by_fundamental(kp1, kp2, matches, 3.0)
    # the keypoints carry the float32 coordinates unchanged, so this is the same call the method made
    F, _ = cv2.findFundamentalMat(p1, p2, cv2.FM_RANSAC, 3.0, 0.99)
    return F, mask.astype(np.uint8)


def main():
    rng = np.random.default_rng(20261105)
    out = {"cv2_version": np.array(cv2.__version__), "K": K, "n_scenes": np.array(N_SCENES)}
    for s in range(N_SCENES):
        n = int(rng.integers(40, 1500))
        p1, p2 = scene(rng, n, float(rng.uniform(0.3, 0.95)), float(rng.uniform(0.1, 1.0)))
        F, fmask = fundamental_through_reference(p1, p2)
        out["s%02d_p1" % s], out["s%02d_p2" % s] = p1, p2
        out["s%02d_F" % s], out["s%02d_Fmask" % s] = F, fmask
        for thr in (1, 3):
            E, emask = ref_utils.calculate_essential_matrix(p1, p2, K, threshold=float(thr))
            out["s%02d_E_%d" % (s, thr)], out["s%02d_Emask_%d" % (s, thr)] = E, emask.ravel().astype(np.uint8)
        print(s, n, int(fmask.sum()), int(out["s%02d_Emask_1" % s].sum()), int(out["s%02d_Emask_3" % s].sum()))
    # small sets: cv2.findFundamentalMat(FM_RANSAC) runs LMedS below 15 points
    for j, n in enumerate((8, 9, 11, 13, 14, 14)):
        p1 = rng.uniform(0, 640, (n, 2)).astype(np.float32)
        p2 = (p1 * 1.01 + rng.normal(0, 0.3, (n, 2)) + np.array([3, 1])).astype(np.float32)
        F, fmask = fundamental_through_reference(p1, p2)
        out["small%d_p1" % j], out["small%d_p2" % j] = p1, p2
        out["small%d_F" % j] = F if F is not None else np.zeros((0, 3))
        out["small%d_Fmask" % j] = fmask
    out["n_small"] = np.array(6)
    path = os.path.join(HERE, "ransac_cases.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
