#!/usr/bin/env python3
"""Generate the golden vectors under tests/golden/ by RUNNING THE REFERENCE ITSELF.

Runs only in the build container (needs /root/reference and cv2); the vectors it writes
are committed so that the tests can run on the GPU box, where the reference is absent.
The reference's hot path is pure Python over OpenCV, so every vector here is what
`/root/reference/src/orbslam2/{extractor,matcher,initializer,utils}.py` return on the
given input with opencv-python-headless (version recorded in each file).

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py
"""
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
import yaml  # noqa: E402

from orbslam2.extractor import ORBExtractor  # noqa: E402  (the reference's)
from orbslam2.initializer import MapInitializer  # noqa: E402
from orbslam2.matcher import DescriptorMatcher  # noqa: E402
from orbslam2 import utils as ref_utils  # noqa: E402
import synth  # noqa: E402  (visual-slam_b200/orbslam2/synth.py, numpy only)

VER = np.array(cv2.__version__)
K_GERARD = np.array([796.7, 0.0, 239.0, 0.0, 787.8, 425.0, 0.0, 0.0, 1.0]).reshape(3, 3)


def kp_arrays(kps):
    return dict(x=np.array([k.pt[0] for k in kps], np.float32), y=np.array([k.pt[1] for k in kps], np.float32),
                size=np.array([k.size for k in kps], np.float32), angle=np.array([k.angle for k in kps], np.float32),
                response=np.array([k.response for k in kps], np.float32),
                octave=np.array([k.octave for k in kps], np.int32))


def match_array(ms):
    return np.array([(m.queryIdx, m.trainIdx, m.distance) for m in ms], np.float64).reshape(-1, 3)


def pref(d, p):
    return {p + k: v for k, v in d.items()}


def extraction_and_matching(tag, img1, img2, n_features, out):
    ex = ORBExtractor(n_features=n_features)
    k1, d1 = ex.detect_and_compute(img1)
    k2, d2 = ex.detect_and_compute(img2)
    out.update(pref(kp_arrays(k1), tag + "_kp1_"))
    out.update(pref(kp_arrays(k2), tag + "_kp2_"))
    out[tag + "_desc1"], out[tag + "_desc2"] = d1, d2
    for r in (0.75, 0.8, 0.85):
        m = DescriptorMatcher(ratio_threshold=r)
        out["%s_match_%d" % (tag, int(r * 100))] = match_array(m.match(d1, d2))
    m = DescriptorMatcher(ratio_threshold=0.75)
    out[tag + "_match_noratio"] = match_array(m.match(d1, d2, ratio_test=False))
    return k1, d1, k2, d2


def geometry(tag, k1, d1, k2, d2, K, out):
    m = DescriptorMatcher(ratio_threshold=0.75)
    matches = m.match(d1, d2)
    p1 = np.float32([k1[x.queryIdx].pt for x in matches])
    p2 = np.float32([k2[x.trainIdx].pt for x in matches])
    out[tag + "_p1"], out[tag + "_p2"] = p1, p2
    filt, mask = m.filter_matches_by_fundamental(k1, k2, matches, 3.0)
    F, _ = cv2.findFundamentalMat(p1, p2, cv2.FM_RANSAC, 3.0, 0.99)
    out[tag + "_F"], out[tag + "_Fmask"] = F, mask.astype(np.uint8)
    for thr in (1.0, 3.0):
        E, emask = ref_utils.calculate_essential_matrix(p1, p2, K, threshold=thr)
        out["%s_E_%d" % (tag, int(thr))] = E
        out["%s_Emask_%d" % (tag, int(thr))] = emask.ravel().astype(np.uint8)
        n, R, t, pmask = ref_utils.recover_pose(E, p1, p2, K, emask.copy())
        out["%s_pose_n_%d" % (tag, int(thr))] = np.array(n)
        out["%s_pose_R_%d" % (tag, int(thr))] = R
        out["%s_pose_t_%d" % (tag, int(thr))] = t
        out["%s_pose_mask_%d" % (tag, int(thr))] = pmask.ravel().astype(np.uint8)
        if thr == 3.0:
            P1 = ref_utils.compute_projection_matrix(np.eye(3), np.zeros((3, 1)), K)
            P2 = ref_utils.compute_projection_matrix(R, t, K)
            sel = pmask.ravel().astype(bool)
            X = ref_utils.triangulate_points(p1[sel], p2[sel], P1, P2)
            out[tag + "_tri_P1"], out[tag + "_tri_P2"], out[tag + "_tri_X"] = P1, P2, X
    init = MapInitializer(K, min_matches=50, min_inliers_ratio=0.8)
    init.set_first_frame(k1, d1, np.zeros((8, 8), np.uint8))
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        ok, R, t, pts, vm = init.initialize(k2, d2, DescriptorMatcher(ratio_threshold=0.75), np.zeros((8, 8), np.uint8))
    out[tag + "_init_ok"] = np.array(bool(ok))
    if ok:
        out[tag + "_init_R"], out[tag + "_init_t"] = R, t
        out[tag + "_init_pos"] = np.array([p["position"] for p in pts], np.float32)
        out[tag + "_init_matches"] = match_array(vm)


def distributed(tag, img, n_features, out):
    ex = ORBExtractor(n_features=n_features)
    kps, desc = ex.distribute_keypoints(img)
    out[tag + "_dist_xy"] = np.array([k.pt for k in kps], np.float32).reshape(-1, 2)
    out[tag + "_dist_desc"] = desc if desc is not None else np.zeros((0, 32), np.uint8)


def main():
    # ---- synthetic 640x480 (BASELINE config C2) --------------------------------
    out = {"cv2_version": VER}
    v = synth.SyntheticVideo(640, 480, 0)
    f0, f1 = v.frame(0), v.frame(1)
    out["frame0_sha"] = np.frombuffer(__import__("hashlib").sha256(f0.tobytes()).digest(), np.uint8)
    k1, d1, k2, d2 = extraction_and_matching("s640", f0, f1, 2000, out)
    geometry("s640", k1, d1, k2, d2, np.array([[320.0, 0, 320], [0, 320, 240], [0, 0, 1]]), out)
    distributed("s640", f0, 2000, out)
    np.savez_compressed(os.path.join(HERE, "synth640.npz"), **out)

    # ---- real fixture pairs (478x850 BGR) ---------------------------------------
    for pair in ("pair01", "pair12"):
        d = os.path.join(REF, "data", "groundtruth_matches", pair)
        b1, b2 = cv2.imread(os.path.join(d, "img1.png")), cv2.imread(os.path.join(d, "img2.png"))
        g1, g2 = cv2.cvtColor(b1, cv2.COLOR_BGR2GRAY), cv2.cvtColor(b2, cv2.COLOR_BGR2GRAY)
        out = {"cv2_version": VER, "gray1": g1, "gray2": g2, "bgr_crop": b1[100:196, 50:178].copy(),
               "gray_crop": g1[100:196, 50:178].copy()}
        k1, d1, k2, d2 = extraction_and_matching(pair, b1, b2, 2000, out)   # BGR in: exercises the gray conversion
        geometry(pair, k1, d1, k2, d2, K_GERARD, out)
        if pair == "pair01":
            distributed(pair, g1, 2000, out)
        np.savez_compressed(os.path.join(HERE, pair + ".npz"), **out)

    # ---- gt_test_matcher procedure on all 20 pairs (src/tests/gt_test_matcher.py:59-132) ----
    out = {"cv2_version": VER}
    ex = ORBExtractor()
    m = DescriptorMatcher(ratio_threshold=0.85)
    prec, rec = [], []
    for i in range(1, 21):
        pair = "pair%02d" % i
        d = os.path.join(REF, "data", "groundtruth_matches", pair)
        pts1, pts2, gt = ref_utils.load_groundtruth_matcher(os.path.join(d, "gt.yaml"))
        img1, img2 = cv2.imread(os.path.join(d, "img1.png")), cv2.imread(os.path.join(d, "img2.png"))
        kp1 = [cv2.KeyPoint(float(x), float(y), 31) for x, y in pts1]
        kp2 = [cv2.KeyPoint(float(x), float(y), 31) for x, y in pts2]
        _, d1 = ex.compute(img1, kp1)
        _, d2 = ex.compute(img2, kp2)
        ms = m.match(d1, d2)
        pred = {(x.queryIdx, x.trainIdx) for x in ms}
        gts = set(gt)
        tp = len(pred & gts)
        p = tp / len(pred) if pred else 0.0
        r = tp / len(gts) if gts else 0.0
        prec.append(p)
        rec.append(r)
        out.update({pair + "_pts1": pts1, pair + "_pts2": pts2, pair + "_gt": np.array(gt, np.int32).reshape(-1, 2),
                    pair + "_desc1": d1, pair + "_desc2": d2, pair + "_match85": match_array(ms)})
    out["precision"], out["recall"] = np.array(prec), np.array(rec)
    print("gt_test_matcher: mean precision %.3f recall %.3f" % (np.mean(prec), np.mean(rec)))
    np.savez_compressed(os.path.join(HERE, "gt_pairs.npz"), **out)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
