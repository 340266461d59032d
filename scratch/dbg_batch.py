import sys, numpy as np
sys.path.insert(0, "visual-slam_b200"); sys.path.insert(0, ".")
from orbslam2._native import OrbHandle
from orbslam2.synth import SyntheticVideo
v = SyntheticVideo(640, 480, 0)
frames = v.frames(0, 3)
h = OrbHandle(640, 480, max_batch=4)
kp, desc, cnt = h.detect_and_compute_host(frames)
for trial in range(3):
    kp1, desc1, cnt1 = h.detect_and_compute_host(frames[1])
    n = int(cnt[1])
    print("cnt", cnt, cnt1)
    for f in kp.dtype.names:
        a, b = kp[1][f], kp1[0][f]
        d = np.nonzero(a.view(np.uint32) != b.view(np.uint32))[0]
        print(trial, f, len(d), d[:10], a[d[:5]], b[d[:5]])
