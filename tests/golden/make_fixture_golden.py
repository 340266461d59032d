#!/usr/bin/env python3
"""Golden vectors for ALL 20 fixture pairs and for C3 / C5-size frames, made by RUNNING THE REFERENCE.

Build container only (needs /root/reference and cv2).  Writes
  tests/golden/fixture_images/pairNN_imgK.png   the 40 fixture images as gray PNGs (the reference's own
                                                cvtColor(BGR2GRAY) of data/groundtruth_matches/pairNN/imgK.png)
  tests/golden/fixtures_all.npz                 per image: sha256 of the reference's keypoints and descriptors in
                                                canonical order (octave, y, x) and the keypoint count; per pair: the
                                                reference's ratio-0.75 match list as rows (x1, y1, octave1, x2, y2,
                                                octave2, distance); for synthetic 1920x1080 / 8000 and 3840x2160 /
                                                16000 frames (regenerated on the test box by orbslam2.synth) the same
                                                hashes.
pair01 and pair12 keep their full arrays in pair01.npz / pair12.npz (make_golden.py).

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_fixture_golden.py
"""
import hashlib
import os
import sys

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "src"))
sys.path.insert(0, os.path.join(ROOT, "visual-slam_b200", "orbslam2"))

import cv2  # noqa: E402
import numpy as np  # noqa: E402

from orbslam2.extractor import ORBExtractor  # noqa: E402  (the reference's)
from orbslam2.matcher import DescriptorMatcher  # noqa: E402
import synth  # noqa: E402


def canonical(kps, desc):
    """(order, hash) of a keypoint set: rows sorted by (octave, y, x); the hash covers every cv2.KeyPoint field the
    reference returns and the descriptor bytes."""
    x = np.array([k.pt[0] for k in kps], np.float32)
    y = np.array([k.pt[1] for k in kps], np.float32)
    o = np.array([k.octave for k in kps], np.int32)
    order = np.lexsort((x, y, o))
    cols = [x, y, np.array([k.size for k in kps], np.float32), np.array([k.angle for k in kps], np.float32),
            np.array([k.response for k in kps], np.float32), o]
    h = hashlib.sha256()
    for c in cols:
        h.update(np.ascontiguousarray(c[order]).tobytes())
    h.update(np.ascontiguousarray(desc[order]).tobytes())
    return order, np.frombuffer(h.digest(), np.uint8)


def match_rows(k1, k2, matches):
    return np.array([(k1[m.queryIdx].pt[0], k1[m.queryIdx].pt[1], k1[m.queryIdx].octave,
                      k2[m.trainIdx].pt[0], k2[m.trainIdx].pt[1], k2[m.trainIdx].octave, m.distance) for m in matches],
                    np.float32).reshape(-1, 7)


def main():
    out = {"cv2_version": np.array(cv2.__version__)}
    os.makedirs(os.path.join(HERE, "fixture_images"), exist_ok=True)
    ex = ORBExtractor(n_features=2000)
    mt = DescriptorMatcher(ratio_threshold=0.75)
    for i in range(1, 21):
        pair = "pair%02d" % i
        res = []
        for w in (1, 2):
            bgr = cv2.imread(os.path.join(REF, "data", "groundtruth_matches", pair, "img%d.png" % w))
            gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)
            cv2.imwrite(os.path.join(HERE, "fixture_images", "%s_img%d.png" % (pair, w)), gray,
                        [cv2.IMWRITE_PNG_COMPRESSION, 9])
            kps, desc = ex.detect_and_compute(bgr)           # the reference converts to gray itself (extractor.py:61-62)
            _, hsh = canonical(kps, desc)
            out["%s_img%d_sha" % (pair, w)] = hsh
            out["%s_img%d_n" % (pair, w)] = np.array(len(kps))
            res.append((kps, desc))
        ms = mt.match(res[0][1], res[1][1])
        out[pair + "_match75"] = match_rows(res[0][0], res[1][0], ms)
        print(pair, len(res[0][0]), len(res[1][0]), len(ms))
    for tag, w, h, nf in (("c3", 1920, 1080, 8000), ("c5", 3840, 2160, 16000)):
        v = synth.SyntheticVideo(w, h, 0)
        exl = ORBExtractor(n_features=nf)
        res = []
        for k in (0, 1):
            f = v.frame(k)
            out["%s_frame%d_sha" % (tag, k)] = np.frombuffer(hashlib.sha256(f.tobytes()).digest(), np.uint8)
            kps, desc = exl.detect_and_compute(f)
            _, hsh = canonical(kps, desc)
            out["%s_kp%d_sha" % (tag, k)] = hsh
            out["%s_kp%d_n" % (tag, k)] = np.array(len(kps))
            res.append((kps, desc))
        ms = mt.match(res[0][1], res[1][1])
        out[tag + "_match75"] = match_rows(res[0][0], res[1][0], ms)
        print(tag, len(res[0][0]), len(res[1][0]), len(ms))
    path = os.path.join(HERE, "fixtures_all.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
