import sys, numpy as np
sys.path.insert(0, "visual-slam_b200"); sys.path.insert(0, ".")
from orbslam2.synth import SyntheticVideo
def stats(img, thr=7, name=""):
    img = img.astype(np.int16)
    H, W = img.shape
    c = img[3:H-3, 3:W-3]
    comp = np.stack([img[6:H, 3:W-3], img[0:H-6, 3:W-3], img[3:H-3, 6:W], img[3:H-3, 0:W-6]])  # (0,+3),(0,-3),(+3,0),(-3,0)
    s = np.sort(comp, axis=0)
    second_largest, second_smallest = s[2], s[1]
    poss = (second_largest > c + thr) | (second_smallest < c - thr)
    full = np.zeros((H, W), bool); full[3:H-3, 3:W-3] = poss
    # warp item = 64 px wide x 2 rows (aligned)
    Hh, Ww = (H // 2) * 2, (W // 64) * 64
    blk = full[:Hh, :Ww].reshape(Hh // 2, 2, Ww // 64, 64).any(axis=(1, 3))
    # thread item = 4 px
    t4 = full[:, :(W // 4) * 4].reshape(H, W // 4, 4).any(axis=2)
    print("%-12s px possible %.3f  4px-items %.3f  warp-items (64x2) needing work %.3f" % (name, poss.mean(), t4.mean(), blk.mean()))
v = SyntheticVideo(640, 480, 0)
stats(v.frame(0), name="synth640")
g = np.load("tests/golden/pair01.npz")
stats(g["gray1"], name="pair01 real")
import cv2
im = g["gray1"]
for l in range(1, 4):
    im = cv2.resize(im, (int(round(im.shape[1] / 1.2)), int(round(im.shape[0] / 1.2))), interpolation=cv2.INTER_LINEAR_EXACT)
    stats(im, name="pair01 L%d" % l)
