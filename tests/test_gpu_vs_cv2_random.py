"""Random cases against cv2 itself (the reference's whole arithmetic is OpenCV; cv2 is in the image of the GPU box).
The golden tests pin the kernels to what the reference produced on its fixtures; these run the same comparisons on
inputs nobody picked: distribute_keypoints (K8), recoverPose (K12), triangulatePoints, undistort, and the F / E
estimators on a few scenes (tools/ransac_agreement_gpu.py runs hundreds)."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
pytestmark = pytest.mark.gpu

K = np.array([[500.0, 0, 320], [0, 500, 240], [0, 0, 1]])


def _scene(rng, n, inlier_ratio=0.8, noise=0.4):
    R, _ = cv2.Rodrigues(rng.normal(0, 0.08, 3))
    t = rng.normal(0, 0.3, 3)
    X = rng.uniform(-2, 2, (n, 3)) + np.array([0, 0, 6.0])
    x1 = (K @ X.T).T
    x2 = (K @ ((R @ X.T).T + t).T).T
    p1 = x1[:, :2] / x1[:, 2:] + rng.normal(0, noise, (n, 2))
    p2 = x2[:, :2] / x2[:, 2:] + rng.normal(0, noise, (n, 2))
    out = rng.random(n) > inlier_ratio
    p2[out] = rng.uniform(0, 640, (int(out.sum()), 2))
    return p1.astype(np.float32), p2.astype(np.float32)


def _texture(rng, w, h):
    img = cv2.GaussianBlur(rng.normal(0, 1, (h, w)).astype(np.float32), (0, 0), float(rng.uniform(1.0, 3.0)))
    img = (img - img.min()) / (img.max() - img.min()) * 200 + 20
    for _ in range(40):
        x, y = int(rng.integers(0, w)), int(rng.integers(0, h))
        cv2.rectangle(img, (x, y), (x + int(rng.integers(5, 60)), y + int(rng.integers(5, 60))), float(rng.uniform(0, 255)), -1)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("seed", range(6))
def test_estimators_pose_and_triangulation_equal_cv2(seed):
    from orbslam2 import utils as U
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(60, 1500))
    p1, p2 = _scene(rng, n, float(rng.uniform(0.4, 0.95)), float(rng.uniform(0.1, 0.8)))
    # F and E: the inlier masks are OpenCV's (same sample stream, same adaptive stop)
    _, mc = cv2.findFundamentalMat(p1, p2, cv2.FM_RANSAC, 3.0, 0.99)
    _, mp = U.find_fundamental_ransac(p1, p2, 3.0, 0.99)
    assert np.array_equal(mp.ravel() != 0, mc.ravel() != 0)
    Ec, mc = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0)
    Ep, mp = U.calculate_essential_matrix(p1, p2, K, cv2.RANSAC, 0.999, 1.0)
    assert np.array_equal(mp.ravel() != 0, mc.ravel() != 0)
    # recoverPose on cv2's own E and mask: same winner, same count, same cheirality mask
    good_c, Rc, tc, pm_c = cv2.recoverPose(Ec, p1, p2, K, mask=mc.copy())
    good_p, Rp, tp, pm_p = U.recover_pose(Ec, p1, p2, K, mask=mc)
    assert good_p == good_c
    assert np.array_equal(pm_p.ravel() != 0, pm_c.ravel() != 0)
    assert np.allclose(Rp, Rc, atol=1e-6) and np.allclose(tp, tc, atol=1e-6)
    # triangulation of the inliers with the recovered pose
    P1 = K @ np.hstack([np.eye(3), np.zeros((3, 1))])
    P2 = K @ np.hstack([Rc, tc])
    inl = pm_c.ravel() != 0
    if inl.sum() >= 2:
        Xc = cv2.triangulatePoints(P1, P2, p1[inl].T.copy(), p2[inl].T.copy())
        Xp = U.triangulate_points(p1[inl], p2[inl], P1, P2)
        xc, xp = Xc[:3] / Xc[3], Xp[:3] / Xp[3]
        assert np.allclose(xp, xc, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("w,h,nf,seed", [(640, 480, 2000, 0), (752, 480, 1000, 1), (1241, 376, 2000, 2), (511, 333, 700, 3)])
def test_distribute_keypoints_equals_the_reference_loop_over_cv2(w, h, nf, seed):
    """extractor.py:85-144: 64 masked goodFeaturesToTrack calls, then ORB.compute at the corners."""
    from orbslam2.extractor import ORBExtractor
    rng = np.random.default_rng(seed)
    img = _texture(rng, w, h)
    orb = cv2.ORB_create(nfeatures=nf)
    ch, cw = h // 8, w // 8
    ref_kps = []
    for i in range(8):
        for j in range(8):
            mask = np.zeros_like(img)
            mask[i * ch:(i + 1) * ch, j * cw:(j + 1) * cw] = 255
            c = cv2.goodFeaturesToTrack(img, maxCorners=nf // 64, qualityLevel=0.01, minDistance=10, mask=mask)
            if c is not None:
                ref_kps += [cv2.KeyPoint(float(x), float(y), 31) for x, y in c.reshape(-1, 2)]
    ref_kps2, ref_desc = orb.compute(img, ref_kps)
    kps, desc = ORBExtractor(n_features=nf).distribute_keypoints(img)
    assert [k.pt for k in kps] == [k.pt for k in ref_kps]          # same corners in the same order
    kps_a, desc_a = ORBExtractor(n_features=nf).distribute_keypoints(img, aligned=True)
    assert [k.pt for k in kps_a] == [k.pt for k in ref_kps2]
    assert np.array_equal(desc_a, ref_desc) and np.array_equal(desc, ref_desc)


@pytest.mark.parametrize("seed", range(4))
def test_undistort_equals_cv2_on_random_cameras(seed):
    from orbslam2 import utils as U
    rng = np.random.default_rng(50 + seed)
    w, h = [(640, 480), (752, 480), (1241, 376), (333, 257)][seed]
    Kc = np.array([[rng.uniform(300, 700), 0, w / 2 + rng.uniform(-20, 20)], [0, rng.uniform(300, 700), h / 2 + rng.uniform(-20, 20)], [0, 0, 1]])
    dist = np.array([rng.uniform(-0.3, 0.1), rng.uniform(-0.05, 0.1), rng.uniform(-1e-3, 1e-3), rng.uniform(-1e-3, 1e-3), rng.uniform(-0.02, 0.02)])
    gray = _texture(rng, w, h)
    assert np.array_equal(U.undistort_image(gray, Kc, dist), cv2.undistort(gray, Kc, dist))
    bgr = np.stack([gray, np.roll(gray, 3, 0), np.roll(gray, 5, 1)], 2)
    assert np.array_equal(U.undistort_image(bgr, Kc, dist), cv2.undistort(bgr, Kc, dist))
