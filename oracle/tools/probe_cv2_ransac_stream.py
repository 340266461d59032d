#!/usr/bin/env python3
"""Probe: is cv2's RANSAC sample stream reproducible from its published algorithm?

cv2.findFundamentalMat(FM_RANSAC) / findEssentialMat(RANSAC) (matcher.py:192, utils.py:124-125,
tracker.py:242 of the reference) are deterministic.  OpenCV's RANSACPointSetRegistrator seeds
`RNG rng((uint64)-1)` per call, draws `idx_i = rng.uniform(0, count)` (multiply-with-carry,
next() % count) with duplicate rejection, and for F rejects a subset whose last point is collinear
with two earlier ones (checkSubset).  The F it returns is the un-refit winning minimal solution,
so the 7 points of the winning sample have ~zero residual: if those 7 indices are a row of the table
generated here, the stream is right.

Run in the build container (uses the committed golden vectors + live cv2 for extra cases).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import geom_oracle as G  # noqa: E402


def main():
    for name, tag in (("synth640", "s640"), ("pair01", "pair01"), ("pair12", "pair12")):
        g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
        p1, p2, F = g[tag + "_p1"], g[tag + "_p2"], g[tag + "_F"]
        n = len(p1)
        x1 = np.c_[p1.astype(np.float64), np.ones(n)]
        x2 = np.c_[p2.astype(np.float64), np.ones(n)]
        r = np.abs(np.einsum("ni,ij,nj->n", x2, F, x1))
        seven = np.sort(np.argsort(r)[:7])
        table = G.cv2_sample_table(p1, p2, 7, 1000, check_collinear=True)
        rows = [i for i, s in enumerate(table) if sorted(s) == seven.tolist()]
        print(name, "n =", n, "winning 7:", seven.tolist(), "residuals <", r[seven].max(),
              "-> table row(s):", rows)
        Fm, mask, info = G.ransac_fundamental_cv2(p1, p2, 3.0, 0.99)
        print("   replay: winner iter", info["best_iter"], "niters", info["niters"], "mask equal:",
              np.array_equal(mask, g[tag + "_Fmask"].astype(bool)), "inliers", int(mask.sum()),
              "max|F - F_cv2| =", np.abs(Fm / Fm[2, 2] - F).max())


if __name__ == "__main__":
    main()
