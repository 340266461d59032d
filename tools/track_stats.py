"""RANSAC horizon of the tracking step on the bench clip: gated matches, inlier ratio and the iteration count the
adaptive stop implies per pair (GPU box).  usage: python tools/track_stats.py"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))
import numpy as np
from orbslam2 import _native, frontend
from orbslam2.synth import SyntheticVideo
B = 256
v = SyntheticVideo(640, 480, 0)
frames = v.frames(0, 64); frames = np.concatenate([frames] * 4)[:B].copy()
h = _native.OrbHandle(640, 480, n_features=2000, max_batch=B)
K = np.array([[320.0, 0, 320], [0, 320, 240], [0, 0, 1]])
out = frontend.track_batched(h, frames, K)
n = out["track_count"].astype(float); inl = out["status"][:, 2].astype(float)
w = inl / np.maximum(n, 1)
it = np.log(1 - 0.999) / np.log(np.maximum(1 - w ** 5, 1e-12))
print("gated matches: median %d  inlier ratio: median %.3f min %.3f  niters: median %.1f p90 %.1f max %.1f  >8: %.0f%%  >40: %.0f%%" % (np.median(n), np.median(w), w.min(), np.median(it), np.percentile(it, 90), it.max(), 100 * (it > 8).mean(), 100 * (it > 40).mean()))
