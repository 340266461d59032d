"""GPU parity of the array-level front end against the reference's own sequence.

N1: `frontend.track_pair` must make the same decisions as tracker.py:214-254 run with cv2 objects: the gated
match list (integer / comparison work) is compared exactly; the essential-matrix stage draws OpenCV's own sample
stream and replays its adaptive stop, so its inlier mask equals cv2.findEssentialMat's on the same point lists.
N2: the BFMatcher stand-in returns exactly cv2.BFMatcher.knnMatch's tuples; the adaptor reroutes a module's
cv2.BFMatcher / cv2.findFundamentalMat.
"""
import math
import types

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

K_SYNTH = np.array([[320.0, 0, 320], [0, 320, 240], [0, 0, 1]])


def _reference_gates(cv2, kps1, kps2, d1, d2, shape, ratio=0.75, pct=0.02):
    """tracker.py:214-230 / matcher.py:70-81,109-169 with cv2 objects (the reference's data flow)."""
    knn = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False).knnMatch(d1, d2, k=2)
    matches = []
    for pair in knn:
        if len(pair) == 2:
            m, n = pair
            if m.distance < ratio * n.distance:
                matches.append(m)
        elif len(pair) == 1:
            matches.append(pair[0])
    h, w = shape[:2]
    max_dist = ((w + h) / 2.0) * pct
    kept = []
    for m in matches:
        a, b = kps1[m.queryIdx].pt, kps2[m.trainIdx].pt
        if math.hypot(b[0] - a[0], b[1] - a[1]) <= max_dist:
            kept.append(m)
    if not kept:
        return []
    kept = sorted(kept, key=lambda x: x.distance)
    thr = np.median([m.distance for m in kept]) * 2.0
    return [m for m in kept if m.distance < thr]


@pytest.mark.parametrize("f0,f1", [(0, 1), (3, 5)])
def test_track_pair_equals_reference_sequence(synth640, f0, f1):
    import cv2
    from orbslam2 import frontend
    from orbslam2.extractor import ORBExtractor
    ex = ORBExtractor(n_features=2000)
    img1, img2 = synth640.frame(f0), synth640.frame(f1)
    kps1, d1 = ex.detect_and_compute(img1)
    kps2, d2 = ex.detect_and_compute(img2)
    pts1 = np.float32([k.pt for k in kps1])
    pts2 = np.float32([k.pt for k in kps2])
    ref = _reference_gates(cv2, kps1, kps2, d1, d2, img1.shape)
    out = frontend.track_pair(pts1, d1, pts2, d2, K_SYNTH, img1.shape)
    assert len(ref) >= 8
    assert out["query_idx"].tolist() == [m.queryIdx for m in ref]
    assert out["train_idx"].tolist() == [m.trainIdx for m in ref]
    assert out["distance"].tolist() == [m.distance for m in ref]
    # essential matrix + pose against cv2 on the same point lists
    p1 = np.float32([kps1[m.queryIdx].pt for m in ref])
    p2 = np.float32([kps2[m.trainIdx].pt for m in ref])
    E, mask = cv2.findEssentialMat(p1, p2, K_SYNTH, cv2.RANSAC, 0.999, 1.0)
    _, R, t, pmask = cv2.recoverPose(E, p1, p2, K_SYNTH, mask=mask.copy())
    assert out["success"] and out["T"].shape == (4, 4)
    a, b = out["ransac_mask"] != 0, mask.ravel() != 0
    assert np.array_equal(a, b), (int(a.sum()), int(b.sum()))       # OpenCV's own sample stream: the same inlier set
    assert not (out["inlier_mask"].astype(bool) & ~a).any() and pmask.shape == mask.shape
    # the synthetic clip is a pan over a plane, where E is not unique (cv2's own cheirality mask keeps a few
    # dozen of ~700 points there): the pose is checked for validity (proper rotation, unit translation, T
    # assembled like tracker.py:250-253), not against cv2's choice among the equivalent solutions
    Ro, to = out["R"], out["t"].ravel()
    assert np.allclose(Ro @ Ro.T, np.eye(3), atol=1e-9) and abs(np.linalg.det(Ro) - 1.0) < 1e-9
    assert abs(np.linalg.norm(to) - 1.0) < 1e-9
    assert np.array_equal(out["T"][:3, :3], Ro) and np.array_equal(out["T"][:3, 3], to) and out["T"][3].tolist() == [0, 0, 0, 1]
    assert R.shape == (3, 3) and t.shape == (3, 1)


def test_track_pair_guards():
    from orbslam2 import frontend
    rng = np.random.default_rng(0)
    d = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    pts = rng.uniform(0, 600, (50, 2)).astype(np.float32)
    out = frontend.track_pair(pts, d, pts[:0], d[:0], K_SYNTH, (480, 640))
    assert not out["success"] and len(out["query_idx"]) == 0
    # identical frames: every match has zero displacement and distance 0 -> median 0 -> `<` keeps none
    out = frontend.track_pair(pts, d, pts, d, K_SYNTH, (480, 640))
    assert not out["success"] and len(out["query_idx"]) == 0
    q, t, dist = frontend.match_arrays(d, d)
    assert q.tolist() == list(range(50)) and t.tolist() == list(range(50)) and not dist.any()


@pytest.mark.parametrize("nq,nt", [(300, 280), (40, 1), (17, 2), (5, 0)])
def test_bfmatcher_stand_in_equals_cv2(nq, nt):
    import cv2
    from orbslam2.frontend import BFMatcher
    rng = np.random.default_rng(nq * 1000 + nt)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    if nt > 3:
        t[3] = t[1]                                           # ties -> lowest train index first
        q[0] = t[1]
    ours = BFMatcher(cv2.NORM_HAMMING, crossCheck=False)
    if nt == 0:
        assert all(len(r) == 0 for r in ours.knnMatch(q, t, k=2))
        return
    ref = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False)
    for k in (1, 2):
        a, b = ours.knnMatch(q, t, k=k), ref.knnMatch(q, t, k=k)
        assert len(a) == len(b)
        for ra, rb in zip(a, b):
            assert [(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in ra] == \
                   [(m.queryIdx, m.trainIdx, m.imgIdx, m.distance) for m in rb]
    assert [(m.queryIdx, m.trainIdx, m.distance) for m in ours.match(q, t)] == \
           [(m.queryIdx, m.trainIdx, m.distance) for m in ref.match(q, t)]


def test_local_mapper_adaptor_reroutes_module_cv2(synth640):
    """A module written like local_mapper.py:116-137 (its own BFMatcher + 0.8 ratio + FM_RANSAC) runs on
    the B200 kernels after install_local_mapper_adaptor, with cv2-compatible results."""
    import cv2
    from orbslam2 import frontend
    from orbslam2.extractor import ORBExtractor
    mod = types.ModuleType("fake_local_mapper")
    mod.cv2 = cv2

    def create_points(kf1, kf2):                              # the reference's statements, on module-level cv2
        matcher = mod.cv2.BFMatcher(mod.cv2.NORM_HAMMING, crossCheck=False)
        matches = matcher.knnMatch(kf1["descriptors"], kf2["descriptors"], k=2)
        good = [p[0] for p in matches if len(p) >= 2 and p[0].distance < 0.8 * p[1].distance]
        p1 = np.float32([kf1["keypoints"][m.queryIdx].pt for m in good])
        p2 = np.float32([kf2["keypoints"][m.trainIdx].pt for m in good])
        F, mask = mod.cv2.findFundamentalMat(p1, p2, mod.cv2.FM_RANSAC, 3.0)
        return good, F, mask

    ex = ORBExtractor(n_features=1500)
    kfs = []
    for f in (0, 2):
        kps, desc = ex.detect_and_compute(synth640.frame(f))
        kfs.append({"keypoints": kps, "descriptors": desc})
    good_ref, F_ref, mask_ref = create_points(*kfs)
    proxy = frontend.install_local_mapper_adaptor(mod)
    assert isinstance(mod.cv2.BFMatcher(cv2.NORM_HAMMING), frontend.BFMatcher) and mod.cv2.FM_RANSAC == cv2.FM_RANSAC
    good, F, mask = create_points(*kfs)
    mod.cv2 = proxy._cv2
    assert [(m.queryIdx, m.trainIdx, m.distance) for m in good] == [(m.queryIdx, m.trainIdx, m.distance) for m in good_ref]
    assert F.shape == (3, 3) and mask.shape == mask_ref.shape and mask.dtype == np.uint8
    assert np.array_equal(mask.ravel() != 0, mask_ref.ravel() != 0)
    assert np.abs(F - F_ref).max() <= 1e-8 * max(1.0, np.abs(F_ref).max())


def test_front_end_full_batched_ransac_equals_per_pair_calls(synth640):
    """BASELINE config 3 (extract + match + RANSAC-F) in one device-resident call: every pair's F, mask and
    inlier count equal vslam_ransac_fundamental_host on that pair's point lists bit for bit (same kernel, same
    seed), the match lists equal the unfused frontend, and the inlier sets meet the cv2 bar of
    tests/test_gpu_geometry.py."""
    import cv2
    from orbslam2 import frontend, utils
    from orbslam2._native import OrbHandle
    frames = synth640.frames(0, 6)
    h = OrbHandle(640, 480, n_features=1000, max_batch=6)
    out = frontend.front_end_full(h, frames, 0.75, 3.0, 0.99)
    kp, desc, cnt, match, mc = h.frontend_host(frames, 0.75)
    assert np.array_equal(out["count"], cnt) and np.array_equal(out["match_count"], mc[:5])
    for p in range(5):
        n = int(mc[p])
        assert n >= 8 and np.array_equal(out["match"][p][:n], match[p][:n])
        q, t = match[p][:n, 0], match[p][:n, 1]
        p1 = np.stack([kp[p]["x"][q], kp[p]["y"][q]], axis=1).astype(np.float32)
        p2 = np.stack([kp[p + 1]["x"][t], kp[p + 1]["y"][t]], axis=1).astype(np.float32)
        F, mask = utils.find_fundamental_ransac(p1, p2, 3.0, 0.99)
        assert np.array_equal(out["F"][p], F) and np.array_equal(out["mask"][p][:n], mask.ravel())
        assert int(out["n_inliers"][p]) == int(mask.sum())
        Fc, mref = cv2.findFundamentalMat(p1, p2, cv2.FM_RANSAC, 3.0, 0.99)
        assert np.array_equal(mask.ravel() != 0, mref.ravel() != 0)
        assert np.abs(F - Fc).max() <= 1e-8 * max(1.0, np.abs(Fc).max())
    h.close()
    # fewer than 8 matches: matcher.py:184-185 -> all ones, no estimate; no matches at all: nothing set
    rng = np.random.default_rng(5)
    noise = rng.integers(0, 256, (2, 240, 320), dtype=np.uint8)
    noise[1] = 255 - noise[1]
    h = OrbHandle(320, 240, n_features=300, max_batch=2)
    out = frontend.front_end_full(h, noise, 0.75, 3.0, 0.99)
    n = int(out["match_count"][0])
    if n < 8:
        assert int(out["n_inliers"][0]) == -1 and out["mask"][0][:n].all() and not out["F"][0].any()
    h.close()


def test_track_batched_equals_track_pair_and_reference_gates(synth640):
    """SURVEY.md 8(f) N1 as one device pipeline: vslam_track_batched (both gates, 5-point RANSAC, recoverPose;
    one CTA per frame pair) gives, for every pair, exactly what the array-level `track_pair` gives -- the
    gated match list is also compared with the reference's cv2-object sequence (tracker.py:214-230), E / masks /
    R / t with the host calls on the same point lists bit for bit (same sampler, solver, error function)."""
    import cv2
    from orbslam2 import frontend
    from orbslam2._native import OrbHandle
    frames = synth640.frames(0, 6)
    h = OrbHandle(640, 480, n_features=2000, max_batch=6)
    out = frontend.track_batched(h, frames, K_SYNTH)
    kp, desc, cnt = out["kp"], out["desc"], out["count"]
    assert out["success"].all() and not out["status"][:, 3].any()
    for p in range(5):
        n0, n1 = int(cnt[p]), int(cnt[p + 1])
        pts0 = np.stack([kp[p]["x"][:n0], kp[p]["y"][:n0]], 1).astype(np.float32)
        pts1 = np.stack([kp[p + 1]["x"][:n1], kp[p + 1]["y"][:n1]], 1).astype(np.float32)
        ref = frontend.track_pair(pts0, desc[p, :n0], pts1, desc[p + 1, :n1], K_SYNTH, frames[0].shape)
        m = int(out["track_count"][p])
        assert m == len(ref["query_idx"]) and m >= 8
        assert np.array_equal(out["track_match"][p, :m, 0], ref["query_idx"])
        assert np.array_equal(out["track_match"][p, :m, 1], ref["train_idx"])
        assert np.array_equal(out["track_match"][p, :m, 2], ref["distance"].astype(np.int32))
        assert ref["success"]
        assert np.array_equal(out["E"][p], ref["E"])
        assert np.array_equal(out["ransac_mask"][p, :m], ref["ransac_mask"])
        assert np.array_equal(out["inlier_mask"][p, :m], ref["inlier_mask"])
        assert np.array_equal(out["R"][p], ref["R"]) and np.array_equal(out["t"][p], ref["t"])
        assert np.array_equal(out["T"][p], ref["T"])
        assert int(out["status"][p, 2]) == int(ref["ransac_mask"].sum())
        assert int(out["status"][p, 1]) == int(ref["inlier_mask"].sum())
        if p == 0:      # the reference's own objects for one pair
            kps0 = [cv2.KeyPoint(float(x), float(y), 31) for x, y in pts0]
            kps1 = [cv2.KeyPoint(float(x), float(y), 31) for x, y in pts1]
            gates = _reference_gates(cv2, kps0, kps1, desc[p, :n0], desc[p + 1, :n1], frames[0].shape)
            assert [[g.queryIdx, g.trainIdx, int(g.distance)] for g in gates] == out["track_match"][p, :m].tolist()
    h.close()
    # degenerate batches: identical frames (all distances 0 -> `<` keeps none -> failure), noise (few matches)
    same = np.stack([frames[0], frames[0]])
    h = OrbHandle(640, 480, n_features=500, max_batch=2)
    out = frontend.track_batched(h, same, K_SYNTH)
    assert not out["success"][0] and int(out["track_count"][0]) == 0 and int(out["status"][0, 0]) == 0
    h.close()
    rng = np.random.default_rng(5)
    noise = rng.integers(0, 256, (2, 240, 320), dtype=np.uint8)
    noise[1] = 255 - noise[1]
    h = OrbHandle(320, 240, n_features=300, max_batch=2)
    out = frontend.track_batched(h, noise, K_SYNTH)
    n0, n1 = int(out["count"][0]), int(out["count"][1])
    pts0 = np.stack([out["kp"][0]["x"][:n0], out["kp"][0]["y"][:n0]], 1).astype(np.float32)
    pts1 = np.stack([out["kp"][1]["x"][:n1], out["kp"][1]["y"][:n1]], 1).astype(np.float32)
    ref = frontend.track_pair(pts0, out["desc"][0, :n0], pts1, out["desc"][1, :n1], K_SYNTH, (240, 320))
    m = int(out["track_count"][0])
    assert m == len(ref["query_idx"]) and bool(out["success"][0]) == bool(ref["success"])
    assert np.array_equal(out["track_match"][0, :m, 0], ref["query_idx"])
    h.close()


def test_track_clip_host_returns_the_poses_of_track_batched(synth640):
    """vslam_track_clip_host: frames in, 4x4 relative poses out (tracker.py:250-253), equal to the batched
    tracking step pair by pair; failed pairs give the identity."""
    from orbslam2 import frontend
    from orbslam2._native import OrbHandle
    frames = synth640.frames(0, 5)
    h = OrbHandle(640, 480, n_features=2000, max_batch=8)
    ref = frontend.track_batched(h, frames, K_SYNTH)
    T, ok, status, count = frontend.track_clip(h, frames, K_SYNTH)
    assert np.array_equal(count, ref["count"]) and np.array_equal(status, ref["status"]) and ok.all()
    assert np.array_equal(T, ref["T"])
    same = np.stack([frames[0], frames[0], frames[1]])
    T, ok, status, _ = frontend.track_clip(h, same, K_SYNTH)
    assert not ok[0] and np.array_equal(T[0], np.eye(4)) and ok[1] and np.array_equal(T[1], ref["T"][0])
    h.close()


class _TrackerStandIn:
    """Tracker.process_frame / _initialize / _track_from_last_frame restated call for call from the reference
    (src/orbslam2/tracker.py:73-146,148-196,198-266) -- the reference's file cannot travel to the GPU box.  It uses the
    module-level `cv2` of this test module exactly where tracker.py uses its own, so that
    frontend.install_tracker_adaptor(module) can reroute findEssentialMat / recoverPose."""

    def __init__(self, camera_matrix, extractor, matcher, initializer, cv2mod):
        self.camera_matrix, self.extractor, self.matcher, self.initializer = camera_matrix, extractor, matcher, initializer
        self.cv2 = cv2mod
        self.state, self.poses, self.map_points = "NOT_INITIALIZED", [], []
        self.last_frame = self.last_keypoints = self.last_descriptors = None
        self.inliers, self.Es, self.Ts = [], [], []

    def process_frame(self, frame):
        cv2 = self.cv2
        current = frame.copy()
        keypoints, descriptors = self.extractor.extract_features(current)            # tracker.py:87 (distributed=True)
        vis = cv2.cvtColor(current, cv2.COLOR_GRAY2BGR) if current.ndim == 2 else current.copy()
        vis = self.extractor.draw_keypoints(vis, keypoints)                          # tracker.py:95
        T = None
        if self.state == "NOT_INITIALIZED":
            if self.last_frame is None:
                self.initializer.set_first_frame(keypoints, descriptors, current)    # tracker.py:162
            else:
                ok, R, t, pts, _ = self.initializer.initialize(keypoints, descriptors, self.matcher, current)   # :168
                if ok:
                    self.map_points, self.state = pts, "TRACKING"
                    self.poses.append(np.eye(4))
        elif self.state == "TRACKING":
            matches = self.matcher.match(self.last_descriptors, descriptors)         # tracker.py:214
            h, w = current.shape[:2]
            matches = self.matcher.filter_matches_by_geometric_distance(self.last_keypoints, keypoints, matches, 0.02,
                                                                        image_shape=(h, w))
            matches = self.matcher.filter_matches_by_distance(matches)
            if len(matches) >= 8:
                p1 = np.float32([self.last_keypoints[m.queryIdx].pt for m in matches])
                p2 = np.float32([keypoints[m.trainIdx].pt for m in matches])
                E, mask = cv2.findEssentialMat(p1, p2, self.camera_matrix, cv2.RANSAC, 0.999, 1.0)     # tracker.py:242
                if E is not None and E.shape == (3, 3):
                    _, R, t, mask = cv2.recoverPose(E, p1, p2, self.camera_matrix, mask=mask)          # tracker.py:249
                    T = np.eye(4)
                    T[:3, :3], T[:3, 3] = R, t.reshape(3)
                    self.poses.append(self.poses[-1] @ T)
                    self.inliers.append(int(np.count_nonzero(mask)))
                    self.Es.append(np.array(E))
                    self.Ts.append(T)
                    inl = [m for i, m in enumerate(matches) if mask[i]]
                    vis = cv2.drawMatches(self.last_frame, self.last_keypoints, current, keypoints, inl[:200], None,
                                          flags=cv2.DrawMatchesFlags_NOT_DRAW_SINGLE_POINTS)
            if T is None:
                self.state = "LOST"
        self.last_frame, self.last_keypoints, self.last_descriptors = current, keypoints, descriptors
        return T, keypoints, vis


def test_tracker_runs_on_the_drop_in_and_equals_the_cv2_run(synth640):
    """'drop-in' means Tracker runs on the swapped classes.  The stand-in above walks
    extract_features(distributed=True) -> draw_keypoints -> set_first_frame / initialize -> match -> both gates ->
    findEssentialMat -> recoverPose over 6 frames, once with every compute call on the B200 kernels (drop-in classes
    + install_tracker_adaptor) and once with the same classes but cv2's own findEssentialMat / recoverPose: the
    kernels run OpenCV's sample stream, so states, inlier counts and poses have to agree."""
    import cv2
    from orbslam2 import frontend
    from orbslam2.extractor import ORBExtractor
    from orbslam2.initializer import MapInitializer
    from orbslam2.matcher import DescriptorMatcher
    K = K_SYNTH
    mod = types.SimpleNamespace(cv2=cv2)
    proxy = frontend.install_tracker_adaptor(mod)
    runs = []
    for cv2mod in (mod.cv2, cv2):
        tr = _TrackerStandIn(K, ORBExtractor(n_features=2000), DescriptorMatcher(), MapInitializer(K), cv2mod)
        out = [tr.process_frame(synth640.frame(k)) for k in range(6)]
        assert all(isinstance(o[2], np.ndarray) and o[2].size > 0 for o in out)
        runs.append((tr, out))
    mod.cv2 = proxy._cv2
    a, b = runs[0][0], runs[1][0]
    assert a.state == b.state and len(a.poses) == len(b.poses) >= 2
    assert a.inliers == b.inliers and len(a.inliers) >= 3          # same inlier sets frame after frame
    assert [len(o[1]) for o in runs[0][1]] == [len(o[1]) for o in runs[1][1]]

    def violation(E):            # how far E is from an essential matrix: 2 E E^T E - tr(E E^T) E and det E
        E = E / np.linalg.norm(E)
        EEt = E @ E.T
        return float(np.linalg.norm(np.r_[(2 * EEt @ E - np.trace(EEt) * E).ravel(), np.linalg.det(E)]))
    exact = 0
    for Ta, Tb, Ea, Eb in zip(a.Ts, b.Ts, a.Es, b.Es):
        assert violation(Ea) < 1e-12
        if violation(Eb) < 1e-9:
            # cv2's E is an essential matrix: same E, same pose
            assert np.allclose(Ta, Tb, atol=1e-6)
            exact += 1
        else:
            # cv2 returned an inexact root of its degree-10 polynomial (constraints off by up to 1e-3 on this clip; the
            # distribute_keypoints misalignment of the reference makes these point sets mostly noise): same inliers,
            # pose equal up to that error
            ang = np.degrees(np.arccos(np.clip((np.trace(Ta[:3, :3] @ Tb[:3, :3].T) - 1) / 2, -1, 1)))
            assert ang < 0.5 and float(Ta[:3, 3] @ Tb[:3, 3]) > 0.99
    assert exact >= 2


@pytest.mark.gpu
def test_frontend_run_graph_replay_equals_plain_launches(synth640):
    """vslam_frontend_run on small batches: the first call with an argument set runs launch by launch, the second
    is captured into a CUDA graph, later ones replay it.  Every call must give what the plain path gives (the
    legacy default stream is never captured: that run is the comparison)."""
    import ctypes as C
    import torch
    from orbslam2 import _native
    L = _native.lib()
    frames = synth640.frames(0, 6)
    h, w = frames.shape[1:]
    orb = _native.OrbHandle(w, h, n_features=1000, max_batch=1)
    cap = orb.kp_capacity
    dev = torch.device("cuda", 0)
    d_frames = torch.from_numpy(frames).to(dev)

    def run(stream_ptr, sync):
        kp = torch.zeros((1, cap, 6), dtype=torch.float32, device=dev)
        desc = [torch.zeros((1, cap, 32), dtype=torch.uint8, device=dev) for _ in range(2)]
        count = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(2)]
        match = torch.zeros((1, cap, 3), dtype=torch.int32, device=dev)
        mcount = torch.zeros(1, dtype=torch.int32, device=dev)
        d_one = torch.empty((1, h, w), dtype=torch.uint8, device=dev)
        outs = []
        # every argument set (the two ping-pong buffer roles) comes round six times: plain, captured, replayed x 4
        for rep in range(12):
            i = rep % 6
            cur, prv = rep & 1, (rep & 1) ^ 1
            d_one.copy_(d_frames[i:i + 1])
            sync()
            _native.check(L.vslam_frontend_run(orb.handle, d_one.data_ptr(), 1, w, desc[prv].data_ptr(), count[prv].data_ptr(),
                                               0.75, kp.data_ptr(), desc[cur].data_ptr(), count[cur].data_ptr(),
                                               match.data_ptr(), mcount.data_ptr(), C.c_void_p(stream_ptr)))
            sync()
            n, m = int(count[cur][0]), int(mcount[0])
            outs.append((n, m, kp[0, :n].cpu().numpy().copy(), desc[cur][0, :n].cpu().numpy().copy(),
                         match[0, :m].cpu().numpy().copy()))
        return outs

    plain = run(0, torch.cuda.synchronize)                      # legacy default stream: launch by launch
    side = torch.cuda.Stream()
    graphed = run(side.cuda_stream, torch.cuda.synchronize)     # captured on its second sight, then replayed
    assert len(plain) == len(graphed)
    for a, b in zip(plain, graphed):
        assert a[0] == b[0] and a[1] == b[1] and a[0] > 300
        for x, y in zip(a[2:], b[2:]):
            assert np.array_equal(x, y)
    orb.close()
