#!/usr/bin/env python3
"""Recover the 256x4 rBRIEF sampling pattern used by cv2.ORB (WTA_K=2, patchSize=31).

TEST/ORACLE TOOLING -- not part of the product path.

The reference (src/orbslam2/extractor.py:38-48,83) delegates descriptor
computation to cv2.ORB, whose sampling table is compiled into the OpenCV wheel
and is not present on disk.  It is recovered by impulse probing (SURVEY.md
section 8(c)): a single bright pixel on a black image makes exactly the tests
whose *second* sample lands on it fire (t0 < t1); a single dark pixel on a
white image makes the tests whose *first* sample lands on it fire.

Writes the table as a C initialiser list to the paths given on the command
line (default: oracle/rbrief_pattern.inc and the product copy under
visual-slam_b200/csrc/), and checks the sha256 recorded in SURVEY.md section 0.
"""
import hashlib
import os
import sys

import cv2
import numpy as np

EXPECTED_SHA256 = "7e645581387b82784797e8adddb9b6f0c12611859fda09ca8a9bec96d767a05f"
R = 18  # probe radius; pattern coordinates lie in [-13, 12]
C = 50  # keypoint centre in the 101x101 probe image


def probe(background, impulse):
    orb = cv2.ORB_create(nfeatures=10, scaleFactor=1.2, nlevels=1, edgeThreshold=31,
                         firstLevel=0, WTA_K=2, scoreType=cv2.ORB_HARRIS_SCORE,
                         patchSize=31, fastThreshold=7)
    found = {}
    for dy in range(-R, R + 1):
        for dx in range(-R, R + 1):
            img = np.full((101, 101), background, np.uint8)
            img[C + dy, C + dx] = impulse
            kp = cv2.KeyPoint(float(C), float(C), 31, 0.0, 0.0, 0)
            _, d = orb.compute(img, [kp])
            bits = np.unpackbits(d[0], bitorder="little")
            for k in np.nonzero(bits)[0]:
                found.setdefault(int(k), []).append((dx, dy))
    return found


def main():
    second = probe(0, 12)          # t1 sample positions (bit fires when t0 < t1)
    first = probe(255, 255 - 12)   # t0 sample positions
    pat = np.zeros((256, 4), np.int32)
    for k in range(256):
        assert len(first.get(k, [])) == 1 and len(second.get(k, [])) == 1, (k, first.get(k), second.get(k))
        pat[k, 0:2] = first[k][0]
        pat[k, 2:4] = second[k][0]
    sha = hashlib.sha256(pat.astype("<i4").tobytes()).hexdigest()
    print("sha256", sha, "sum", int(pat.sum()), "abs-sum", int(np.abs(pat).sum()),
          "range", int(pat.min()), int(pat.max()))
    assert sha == EXPECTED_SHA256, "pattern hash differs from SURVEY.md section 0"
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(os.path.dirname(here))
    outs = sys.argv[1:] or [os.path.join(root, "oracle", "rbrief_pattern.inc"),
                            os.path.join(root, "visual-slam_b200", "csrc", "rbrief_pattern.inc")]
    body = ["/* rBRIEF 256x4 sampling pattern {x0,y0,x1,y1} of cv2.ORB (WTA_K=2, patchSize=31),",
            " * recovered from opencv-python-headless %s by oracle/tools/recover_pattern.py." % cv2.__version__,
            " * sha256(int32 LE [256][4]) = %s */" % sha]
    for k in range(256):
        body.append("{%d,%d,%d,%d}," % tuple(int(v) for v in pat[k]))
    for o in outs:
        with open(o, "w") as f:
            f.write("\n".join(body) + "\n")
        print("wrote", o)


if __name__ == "__main__":
    main()
