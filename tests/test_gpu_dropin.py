"""GPU tests of the drop-in classes (same names / signatures / return types as the reference's
src/orbslam2 modules) against golden vectors produced by the reference's own classes."""
import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")

K_GERARD = np.array([796.7, 0.0, 239.0, 0.0, 787.8, 425.0, 0.0, 0.0, 1.0]).reshape(3, 3)


def _kp_set(kps, desc):
    return {(k.octave, k.pt[0], k.pt[1]): (k.angle, k.response, k.size, bytes(d)) for k, d in zip(kps, desc)}


def _golden_set(g, tag, which):
    pre = "%s_kp%d_" % (tag, which)
    return {(int(o), float(x), float(y)): (float(a), float(r), float(s), bytes(d))
            for x, y, o, a, r, s, d in zip(g[pre + "x"], g[pre + "y"], g[pre + "octave"], g[pre + "angle"],
                                           g[pre + "response"], g[pre + "size"], g["%s_desc%d" % (tag, which)])}


def test_detect_and_compute_equals_reference(synth640):
    from orbslam2.extractor import ORBExtractor
    g = load_golden("synth640.npz")
    ex = ORBExtractor(n_features=2000)
    assert (ex.n_features, ex.scale_factor, ex.n_levels, ex.ini_threshold, ex.min_threshold) == (2000, 1.2, 8, 20, 7)
    kps, desc = ex.detect_and_compute(synth640.frame(0))
    assert isinstance(kps, tuple) and isinstance(kps[0], cv2.KeyPoint) and kps[0].class_id == -1
    assert desc.dtype == np.uint8 and desc.shape == (len(kps), 32) and desc.flags["C_CONTIGUOUS"]
    assert _kp_set(kps, desc) == _golden_set(g, "s640", 1)          # 100 % of keypoints, bit-equal fields
    assert [k.octave for k in kps] == sorted(k.octave for k in kps)  # level-major like cv2
    # blank image -> ((), None), and then match() -> []
    kps0, desc0 = ex.detect_and_compute(np.full((480, 640), 90, np.uint8))
    assert kps0 == () and desc0 is None
    from orbslam2.matcher import DescriptorMatcher
    assert DescriptorMatcher().match(desc0, desc) == []


@pytest.mark.parametrize("pair", ["pair01", "pair12"])
def test_real_bgr_images_and_match_lists(pair):
    from orbslam2.extractor import ORBExtractor
    from orbslam2.matcher import DescriptorMatcher
    g = load_golden(pair + ".npz")
    ex = ORBExtractor()
    # K0: BGR input goes through the gray conversion kernel
    crop = ex._gray(g["bgr_crop"])
    assert np.array_equal(crop, g["gray_crop"])
    k1, d1 = ex.detect_and_compute(g["gray1"])
    k2, d2 = ex.detect_and_compute(g["gray2"])
    assert _kp_set(k1, d1) == _golden_set(g, pair, 1)
    assert _kp_set(k2, d2) == _golden_set(g, pair, 2)
    # matching is checked on the reference's descriptor arrays (its in-level keypoint order differs)
    r1, r2 = g[pair + "_desc1"], g[pair + "_desc2"]
    for r in (75, 80, 85):
        ms = DescriptorMatcher(ratio_threshold=r / 100.0).match(r1, r2)
        assert isinstance(ms[0], cv2.DMatch) and ms[0].imgIdx == 0
        got = np.array([(m.queryIdx, m.trainIdx, m.distance) for m in ms], np.float64)
        assert np.array_equal(got, g["%s_match_%d" % (pair, r)])
    ms = DescriptorMatcher().match(r1, r2, ratio_test=False)
    got = np.array([(m.queryIdx, m.trainIdx, m.distance) for m in ms], np.float64)
    assert np.array_equal(got, g[pair + "_match_noratio"])
    # non-uint8 descriptors are cast like the reference does
    ms2 = DescriptorMatcher().match(r1.astype(np.int32), r2.astype(np.float32), ratio_test=False)
    assert len(ms2) == len(ms)


@pytest.mark.parametrize("pair", ["pair%02d" % i for i in range(1, 21)])
def test_every_fixture_image_equals_reference(pair):
    """All 40 fixture images through the drop-in ORBExtractor / DescriptorMatcher: keypoints (every field) and
    descriptors hash to the reference's, the match list equals the reference's as a set of rows (up to a handful of
    equal-distance ties, which cv2 breaks by its unspecified in-level keypoint order)."""
    from conftest import canonical_kp_hash, fixture_image, match_row_set
    from orbslam2.extractor import ORBExtractor
    from orbslam2.matcher import DescriptorMatcher
    g = load_golden("fixtures_all.npz")
    ex = ORBExtractor(n_features=2000)
    res = []
    for w in (1, 2):
        kps, desc = ex.detect_and_compute(fixture_image(pair, w))
        assert len(kps) == int(g["%s_img%d_n" % (pair, w)])
        hsh = canonical_kp_hash([k.pt[0] for k in kps], [k.pt[1] for k in kps], [k.size for k in kps],
                                [k.angle for k in kps], [k.response for k in kps], [k.octave for k in kps], desc)
        assert np.array_equal(hsh, g["%s_img%d_sha" % (pair, w)])
        res.append((kps, desc))
    (k1, d1), (k2, d2) = res
    ms = DescriptorMatcher(ratio_threshold=0.75).match(d1, d2)
    rows = [(k1[m.queryIdx].pt[0], k1[m.queryIdx].pt[1], k1[m.queryIdx].octave, k2[m.trainIdx].pt[0],
             k2[m.trainIdx].pt[1], k2[m.trainIdx].octave, m.distance) for m in ms]
    got, ref = match_row_set(rows), match_row_set(g[pair + "_match75"])
    assert len(got ^ ref) <= 6, (len(got), len(ref), len(got ^ ref))


@pytest.mark.parametrize("tag,w,h,nf", [("c3", 1920, 1080, 8000), ("c5", 3840, 2160, 16000)])
def test_c3_and_c5_size_frames_equal_reference(tag, w, h, nf):
    """Reference-made goldens at the sizes of BASELINE configs 3 and 5."""
    from conftest import canonical_kp_hash, match_row_set
    from orbslam2 import _native
    from orbslam2.synth import SyntheticVideo
    g = load_golden("fixtures_all.npz")
    v = SyntheticVideo(w, h, 0)
    frames = v.frames(0, 2)
    hd = _native.OrbHandle(w, h, n_features=nf, max_batch=2)
    kp, desc, cnt, match, mc = hd.frontend_host(frames, 0.75)
    for k in (0, 1):
        n = int(cnt[k])
        assert n == int(g["%s_kp%d_n" % (tag, k)])
        a = kp[k][:n]
        assert np.array_equal(canonical_kp_hash(a["x"], a["y"], a["size"], a["angle"], a["response"], a["octave"], desc[k][:n]),
                              g["%s_kp%d_sha" % (tag, k)])
    rows = [(kp[0]["x"][q], kp[0]["y"][q], kp[0]["octave"][q], kp[1]["x"][t], kp[1]["y"][t], kp[1]["octave"][t], d)
            for q, t, d in match[0][:mc[0]]]
    got, ref = match_row_set(rows), match_row_set(g[tag + "_match75"])
    assert len(got ^ ref) <= 12, (len(got), len(ref), len(got ^ ref))
    hd.close()


def test_gt_matcher_procedure_and_compute():
    """src/tests/gt_test_matcher.py on the committed pairs: compute() at hand-labelled keypoints
    then match(0.85) -> the reference's exact match list."""
    from orbslam2.extractor import ORBExtractor
    from orbslam2.matcher import DescriptorMatcher
    gt = load_golden("gt_pairs.npz")
    ex, m = ORBExtractor(), DescriptorMatcher(ratio_threshold=0.85)
    for pair in ("pair01", "pair12"):
        g = load_golden(pair + ".npz")
        descs = []
        for which in (1, 2):
            pts = gt["%s_pts%d" % (pair, which)]
            kps = [cv2.KeyPoint(float(x), float(y), 31) for x, y in pts]
            out_kps, d = ex.compute(g["gray%d" % which], kps)
            assert out_kps is kps                                  # the reference returns the caller's list
            assert np.array_equal(d, gt["%s_desc%d" % (pair, which)])
            descs.append(d)
        ms = m.match(descs[0], descs[1])
        got = np.array([(x.queryIdx, x.trainIdx, x.distance) for x in ms], np.float64)
        assert np.array_equal(got, gt[pair + "_match85"])
    # the matcher alone on all 20 pairs (descriptors from the reference)
    for i in range(1, 21):
        p = "pair%02d" % i
        ms = m.match(gt[p + "_desc1"], gt[p + "_desc2"])
        got = np.array([(x.queryIdx, x.trainIdx, x.distance) for x in ms], np.float64).reshape(-1, 3)
        assert np.array_equal(got, gt[p + "_match85"])


def test_compute_border_and_octaves(oracle, synth640):
    from orbslam2.extractor import ORBExtractor
    frame = synth640.frame(2)
    rng = np.random.default_rng(3)
    kps = [cv2.KeyPoint(float(rng.uniform(0, 640)), float(rng.uniform(0, 480)), 31, float(rng.uniform(-1, 360)), 0,
                        int(rng.integers(0, 4))) for _ in range(300)]
    ex = ORBExtractor()
    out, d = ex.compute(frame, kps)
    ref, keep = oracle.orb_compute(frame, [k.pt[0] for k in kps], [k.pt[1] for k in kps], [k.angle for k in kps],
                                   [k.octave for k in kps])
    assert out is kps and len(d) == len(ref) < len(kps)
    assert np.array_equal(d, ref)
    al_kps, al_d = ex.compute(frame, kps, aligned=True)
    assert [kps[i] for i in keep] == al_kps and np.array_equal(al_d, ref)


@pytest.mark.parametrize("name,tag", [("synth640.npz", "s640"), ("pair01.npz", "pair01")])
def test_distribute_keypoints_equals_reference(name, tag, synth640):
    from orbslam2.extractor import ORBExtractor
    g = load_golden(name)
    img = synth640.frame(0) if tag == "s640" else g["gray1"]
    kps, desc = ORBExtractor(n_features=2000).distribute_keypoints(img)
    assert isinstance(kps, list)
    xy = np.array([k.pt for k in kps], np.float32)
    assert np.array_equal(xy, g[tag + "_dist_xy"])                   # same corners, same order
    assert all(k.angle == -1 and k.octave == 0 and k.size == 31 for k in kps)
    assert np.array_equal(desc, g[tag + "_dist_desc"])              # N_all keypoints, N_in-border rows
    assert len(desc) < len(kps)
    k2, d2 = ORBExtractor(n_features=2000).extract_features(img)     # distributed=True is the default
    assert len(k2) == len(kps) and np.array_equal(d2, desc)


def test_match_with_mask(oracle):
    from orbslam2.matcher import DescriptorMatcher
    rng = np.random.default_rng(5)
    d1 = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    d2 = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    mask = (rng.random((40, 300)) < 0.2).astype(np.uint8)
    mask[7] = 0
    ms = DescriptorMatcher().match_with_mask(d1, d2, mask)
    idx, dist = oracle.bf_match_masked(d1, d2, mask)
    exp = sorted([(float(dist[q]), q, int(idx[q])) for q in range(40) if idx[q] >= 0], key=lambda t: t[0])
    assert [(m.distance, m.queryIdx, m.trainIdx) for m in ms] == exp


def test_filter_by_fundamental_and_initializer():
    from orbslam2.extractor import ORBExtractor
    from orbslam2.initializer import MapInitializer
    from orbslam2.matcher import DescriptorMatcher
    g = load_golden("pair01.npz")
    ex, m = ORBExtractor(), DescriptorMatcher()
    k1, d1 = ex.detect_and_compute(g["gray1"])
    k2, d2 = ex.detect_and_compute(g["gray2"])
    matches = m.match(d1, d2)
    assert abs(len(matches) - len(g["pair01_match_75"])) <= 2      # same set up to in-level order ties
    filt, mask = m.filter_matches_by_fundamental(k1, k2, matches, 3.0)
    assert mask.dtype == bool and len(mask) == len(matches) and len(filt) == int(mask.sum())
    assert int(mask.sum()) >= 0.97 * int(g["pair01_Fmask"].sum())   # own in-level keypoint order -> another point order
    init = MapInitializer(K_GERARD, min_matches=50, min_inliers_ratio=0.8)
    assert init.initialize(k2, d2, m, g["gray2"]) == (False, None, None, None, None)
    # From the reference's own keypoints (cv2's in-level order) everything downstream is the reference's result:
    # the match list, the fundamental mask, and the whole initialisation (E-RANSAC on OpenCV's sample stream,
    # recoverPose, triangulation, depth filter) -- same pose, same 213 map points.
    import cv2
    gk = [[cv2.KeyPoint(float(x), float(y), float(sz), float(a), float(r), int(o)) for x, y, sz, a, r, o in
           zip(*(g["pair01_kp%d_%s" % (w, f)] for f in ("x", "y", "size", "angle", "response", "octave")))] for w in (1, 2)]
    gd1, gd2 = g["pair01_desc1"], g["pair01_desc2"]
    gm = m.match(gd1, gd2)
    assert [(x.queryIdx, x.trainIdx, x.distance) for x in gm] == [tuple(r) for r in g["pair01_match_75"].tolist()]
    _, gmask = m.filter_matches_by_fundamental(gk[0], gk[1], gm, 3.0)
    assert np.array_equal(gmask, g["pair01_Fmask"] != 0)
    init.set_first_frame(gk[0], gd1, g["gray1"])
    ok, R, t, pts, vm = init.initialize(gk[1], gd2, m, g["gray2"])
    assert ok is True and R.shape == (3, 3) and t.shape == (3, 1) and init.initialization_done
    assert set(pts[0]) == {"position", "color", "keypoint_references", "observed_frames"}
    assert pts[0]["position"].dtype == np.float32 and pts[0]["position"].shape == (3,)
    assert np.allclose(R, g["pair01_init_R"], atol=1e-6) and np.allclose(t, g["pair01_init_t"], atol=1e-6)
    assert len(pts) == len(g["pair01_init_pos"]) == len(vm)
    assert [(x.queryIdx, x.trainIdx, x.distance) for x in vm] == [tuple(r) for r in g["pair01_init_matches"].tolist()]
    pos = np.array([q["position"] for q in pts])
    assert np.allclose(pos, g["pair01_init_pos"], rtol=1e-3, atol=1e-3)
    vis = init.draw_initialization(g["gray1"], g["gray2"], vm[:50])
    assert isinstance(vis, np.ndarray) and vis.size > 0


def test_bgr_frames_ingested_on_device(synth640):
    """Colour frames (what Tracker.process_frame hands to the extractor, tracker.py:84-87) are uploaded once and
    converted on the device (K0) in front of the pyramid: same keypoints / descriptors as converting first, and
    the conversion equals cv2's fixed-point formula on a sweep of colours incl. the unaligned tail."""
    import cv2
    from orbslam2 import _native
    from orbslam2.extractor import ORBExtractor
    gray = synth640.frame(2)
    rng = np.random.default_rng(3)
    bgr = np.stack([np.clip(gray.astype(np.int16) + rng.integers(-40, 41, gray.shape), 0, 255).astype(np.uint8)
                    for _ in range(3)], axis=2)
    ref_gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)     # type container only elsewhere; here the K0 golden formula
    ex = ORBExtractor(n_features=1500)
    k_bgr, d_bgr = ex.detect_and_compute(bgr)
    k_gray, d_gray = ex.detect_and_compute(ref_gray)
    assert len(k_bgr) == len(k_gray) and np.array_equal(d_bgr, d_gray)
    assert [(k.pt, k.angle, k.response, k.octave) for k in k_bgr] == [(k.pt, k.angle, k.response, k.octave) for k in k_gray]
    # batch of two frames through the handle
    hd = _native.OrbHandle(640, 480, n_features=1500, max_batch=2)
    kp2, desc2, cnt2 = hd.detect_and_compute_host_bgr(np.stack([bgr, bgr[:, ::-1].copy()]))
    kp1, desc1, cnt1 = hd.detect_and_compute_host(np.stack([ref_gray, cv2.cvtColor(bgr[:, ::-1].copy(), cv2.COLOR_BGR2GRAY)]))
    assert np.array_equal(cnt1, cnt2) and np.array_equal(desc1, desc2) and np.array_equal(kp1, kp2)
    hd.close()
    # conversion alone on odd sizes (byte tail) against the formula
    for h, w in ((7, 13), (1, 1), (33, 31)):
        c = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        out = np.empty((h, w), np.uint8)
        _native.check(_native.lib().vslam_bgr2gray_host(_native.ptr(c), h, w, _native.ptr(out)))
        want = ((c[..., 0].astype(np.int64) * 3735 + c[..., 1].astype(np.int64) * 19235 + c[..., 2].astype(np.int64) * 9798 + 16384) >> 15).astype(np.uint8)
        assert np.array_equal(out, want) and np.array_equal(out, cv2.cvtColor(c, cv2.COLOR_BGR2GRAY))
