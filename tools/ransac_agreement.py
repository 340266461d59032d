#!/usr/bin/env python3
"""How often does a replay of OpenCV's RANSAC loop (its RNG stream, its adaptive stop) reproduce what cv2 returns?

Build container only (needs cv2).  For synthetic two-view scenes over a range of inlier ratios and noise, compares
the inlier mask of cv2.findFundamentalMat(FM_RANSAC) / cv2.findEssentialMat(RANSAC) with
  * oracle: oracle/geom_oracle.py (numpy solvers, every real solution polished),
  * product: the same loop around the product's own minimal solvers compiled for the host
    (tools/host_harness/solvers_host.cpp; the CUDA kernels run this very code per thread).
Prints, per estimator, the fraction of scenes with an identical mask and the worst Jaccard.
"""
import ctypes as C
import os
import subprocess
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import geom_oracle as G  # noqa: E402

SO = "/tmp/libvs_solvers_host.so"


def host_solvers():
    src = os.path.join(ROOT, "tools", "host_harness", "solvers_host.cpp")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", SO, src])
    L = C.CDLL(SO)
    dp = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    L.h_solve_5pt.argtypes = [dp, dp, dp]
    L.h_solve_7pt.argtypes = [dp, dp, dp]

    def five(q1, q2):
        out = np.zeros((10, 9))
        n = L.h_solve_5pt(np.ascontiguousarray(q1, np.float64), np.ascontiguousarray(q2, np.float64), out)
        return [out[k].reshape(3, 3) for k in range(n)]

    def seven(m1, m2):
        # the kernels hand the solver Hartley-normalised points; the solution set does not depend on it
        out = np.zeros((3, 9))
        n = L.h_solve_7pt(np.ascontiguousarray(m1, np.float64), np.ascontiguousarray(m2, np.float64), out)
        return [out[k].reshape(3, 3) for k in range(n)]
    return five, seven


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
    n_scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    with_oracle = "--product-only" not in sys.argv
    five, seven = host_solvers()
    K = np.array([[500.0, 0, 320], [0, 500, 240], [0, 0, 1]])
    rng = np.random.default_rng(7)
    stats = {}
    for s in range(n_scenes):
        n = int(rng.integers(40, 1500))
        ratio = float(rng.uniform(0.3, 0.95))
        noise = float(rng.uniform(0.1, 1.0))
        p1, p2 = scene(rng, n, ratio, noise, K)
        Fc, mc = cv2.findFundamentalMat(p1, p2, cv2.FM_RANSAC, 3.0, 0.99)
        mc = mc.ravel().astype(bool)
        _, mp, ip = G._cv2_ransac(p1, p2, 7, seven, G.fundamental_error, 3.0, 0.99, 1000, True)
        if with_oracle:
            _, mo, io = G.ransac_fundamental_cv2(p1, p2, 3.0, 0.99)
            stats.setdefault("F oracle", []).append((np.array_equal(mo, mc), jacc(mo, mc), io["niters"]))
        stats.setdefault("F product-solver", []).append((np.array_equal(mp, mc), jacc(mp, mc), ip["niters"]))
        for thr in (1.0, 3.0):
            Ec, mc = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, thr)
            mc = mc.ravel().astype(bool)
            q1, q2 = G.normalise(p1, K), G.normalise(p2, K)
            t = thr / ((K[0, 0] + K[1, 1]) / 2)
            _, mp, ip = G._cv2_ransac(q1, q2, 5, five, G.essential_error, t, 0.999, 1000, False)
            if with_oracle:
                _, mo, io = G.ransac_essential_cv2(p1, p2, K, 0.999, thr)
                stats.setdefault("E thr %g oracle" % thr, []).append((np.array_equal(mo, mc), jacc(mo, mc), io["niters"]))
            stats.setdefault("E thr %g product-solver" % thr, []).append((np.array_equal(mp, mc), jacc(mp, mc), ip["niters"]))
    for k, v in stats.items():
        eq = np.mean([x[0] for x in v])
        js = np.array([x[1] for x in v])
        print("%-26s identical masks %5.1f %%   Jaccard min %.4f  mean %.4f   median niters %d" %
              (k, 100 * eq, js.min(), js.mean(), int(np.median([x[2] for x in v]))))


if __name__ == "__main__":
    main()
