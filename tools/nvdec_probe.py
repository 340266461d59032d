"""GPU-box probe for the NVDEC ingest: writes a short VP9 clip with cv2, decodes it (i) with cv2 on the CPU and
(ii) with orbslam2.ingest.VideoReader on NVDEC, compares them and dumps (NV12, BGR) pairs for fitting the swscale constants."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200"))
import cv2  # noqa: E402
import numpy as np  # noqa: E402

from orbslam2 import ingest  # noqa: E402
from orbslam2.synth import SyntheticVideo  # noqa: E402

out_dir = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
fourcc, ext = (sys.argv[2], sys.argv[3]) if len(sys.argv) > 3 else ("VP90", "webm")
print("nvdec available:", ingest.available())
v = SyntheticVideo(640, 480, 0)
rng = np.random.default_rng(0)
path = "/tmp/probe.%s" % ext
w = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*fourcc), 30.0, (640, 480))
assert w.isOpened()
for k in range(10):
    g = v.frame(k)
    bgr = cv2.cvtColor(g, cv2.COLOR_GRAY2BGR)
    # colour so that U and V vary: tint blocks of the frame
    tint = rng.integers(-60, 61, (480 // 32, 640 // 32, 3)).astype(np.int16)
    bgr = np.clip(bgr.astype(np.int16) + np.kron(tint, np.ones((32, 32, 1), np.int16)), 0, 255).astype(np.uint8)
    w.write(bgr)
w.release()
cap = cv2.VideoCapture(path, cv2.CAP_FFMPEG)
ref_bgr = []
while True:
    ok, f = cap.read()
    if not ok:
        break
    ref_bgr.append(f)
cap.release()
cap = cv2.VideoCapture(path, cv2.CAP_FFMPEG)
cap.set(cv2.CAP_PROP_CONVERT_RGB, 0)
ref_y = []
while True:
    ok, f = cap.read()
    if not ok:
        break
    ref_y.append(f.copy())
cap.release()
print("cv2 frames:", len(ref_bgr), ref_bgr[0].shape, "raw:", ref_y[0].shape)
rd = ingest.VideoReader(path, mode="nv12")
nv12 = [f.cpu().numpy() for f in rd]
rd.close()
print("nvdec frames:", len(nv12), nv12[0].shape if nv12 else None)
n = min(len(nv12), len(ref_y))
for i in range(n):
    yd = nv12[i][:480]
    print(i, "luma equal:", np.array_equal(yd, ref_y[i][:480, :640]), "maxdiff", int(np.abs(yd.astype(int) - ref_y[i][:480, :640].astype(int)).max()))
np.savez_compressed(os.path.join(out_dir, "nvdec_probe.npz"), nv12=np.stack(nv12[:n]), bgr=np.stack(ref_bgr[:n]))
for mode in ("gray", "bgr"):
    rd = ingest.VideoReader(path, mode=mode)
    fr = [f.cpu().numpy() for f in rd]
    rd.close()
    for i in range(min(3, len(fr))):
        ref = cv2.cvtColor(ref_bgr[i], cv2.COLOR_BGR2GRAY) if mode == "gray" else ref_bgr[i]
        d = np.abs(fr[i].astype(int) - ref.astype(int))
        print(mode, i, "equal:", bool((d == 0).all()), "maxdiff", int(d.max()), "mismatch px", int((d > 0).sum()))
