"""One device-resident tracking step (extract + match + vslam_track_batched) for ncu captures of the RANSAC kernels."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from orbslam2 import _native, frontend  # noqa: E402
from orbslam2.synth import SyntheticVideo  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
v = SyntheticVideo(640, 480, 0)
frames = v.frames(0, min(B, 64))
frames = np.concatenate([frames] * ((B + 63) // 64))[:B].copy()
h = _native.OrbHandle(640, 480, n_features=2000, max_batch=B)
K = np.array([[320.0, 0, 320], [0, 320, 240], [0, 0, 1]])
for _ in range(2):
    out = frontend.track_batched(h, frames, K)
print("pairs ok", int(out["success"].sum()), "of", B - 1)
