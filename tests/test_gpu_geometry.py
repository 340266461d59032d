"""GPU parity of the geometry kernels (C ABI through orbslam2.utils) against the numpy
oracle and the golden vectors produced by the reference (cv2 4.13).

Bars: scoring masks exact given the reference's own F / E; pose: identical winner, mask equal
to the oracle's, R / t within 1e-6; triangulation rel 1e-4 on xyz/w; full estimators
(our own seeded sampler, so a different hypothesis stream than cv2's RNG): inlier-set Jaccard
against cv2's mask no worse than cv2 agrees with ITSELF when its input order is permuted
(measured in the build container over 8 permutations of the same three fixtures:
F thr 3: 0.942-0.968, E thr 1: 0.82-0.86, E thr 3: 0.991) and at least 97 % of cv2's inlier count.
"""
import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu

K_GERARD = np.array([796.7, 0.0, 239.0, 0.0, 787.8, 425.0, 0.0, 0.0, 1.0]).reshape(3, 3)
K_SYNTH = np.array([[320.0, 0, 320], [0, 320, 240], [0, 0, 1]])
CASES = [("synth640.npz", "s640", K_SYNTH), ("pair01.npz", "pair01", K_GERARD), ("pair12.npz", "pair12", K_GERARD)]


def _jaccard(a, b):
    a, b = a.ravel() != 0, b.ravel() != 0
    return (a & b).sum() / max((a | b).sum(), 1)


@pytest.mark.parametrize("name,tag,K", CASES)
def test_scoring_masks_exact(name, tag, K):
    from oracle import geom_oracle as G
    from orbslam2 import utils
    g = load_golden(name)
    p1, p2 = g[tag + "_p1"], g[tag + "_p2"]
    got = utils.score_fundamental(g[tag + "_F"], p1, p2, 3.0)
    assert np.array_equal(got, g[tag + "_Fmask"])                      # cv2's own mask for cv2's own F
    assert np.array_equal(got, G.score_fundamental(g[tag + "_F"], p1, p2, 3.0))
    for thr in (1, 3):
        E = g["%s_E_%d" % (tag, thr)]
        got = utils.score_essential(E, p1, p2, K, float(thr))
        assert np.array_equal(got, g["%s_Emask_%d" % (tag, thr)])
    # a perturbed model: still identical to the oracle point by point
    rng = np.random.default_rng(0)
    F2 = g[tag + "_F"] * (1 + 1e-3 * rng.standard_normal((3, 3)))
    assert np.array_equal(utils.score_fundamental(F2, p1, p2, 1.5), G.score_fundamental(F2, p1, p2, 1.5))


@pytest.mark.parametrize("name,tag,K", CASES)
def test_recover_pose_and_triangulate(name, tag, K):
    from oracle import geom_oracle as G
    from orbslam2 import utils
    g = load_golden(name)
    p1, p2 = g[tag + "_p1"], g[tag + "_p2"]
    for thr in (1, 3):
        E, emask = g["%s_E_%d" % (tag, thr)], g["%s_Emask_%d" % (tag, thr)]
        n, R, t, m = utils.recover_pose(E, p1, p2, K, emask.reshape(-1, 1))
        assert n == int(g["%s_pose_n_%d" % (tag, thr)])
        assert m.shape == (len(p1), 1) and m.dtype == np.uint8
        assert np.array_equal(m.ravel() != 0, g["%s_pose_mask_%d" % (tag, thr)] != 0)
        assert np.allclose(R, g["%s_pose_R_%d" % (tag, thr)], atol=1e-6)
        assert np.allclose(t, g["%s_pose_t_%d" % (tag, thr)], atol=1e-6)
        assert t.shape == (3, 1) and R.dtype == np.float64
        # without a mask the oracle is the reference
        n2, R2, t2, m2 = utils.recover_pose(E, p1, p2, K)
        on, oR, ot, om = G.recover_pose(E, p1, p2, K)
        assert n2 == on and np.array_equal(m2.ravel(), om) and np.allclose(R2, oR, atol=1e-6)
    sel = g[tag + "_pose_mask_3"] != 0
    X = utils.triangulate_points(p1[sel], p2[sel], g[tag + "_tri_P1"], g[tag + "_tri_P2"])
    ref = g[tag + "_tri_X"]
    assert X.shape == ref.shape and X.dtype == np.float32
    a, b = X[:3] / X[3], ref[:3] / ref[3]
    assert np.allclose(a, b, rtol=1e-4, atol=1e-4 * np.abs(b).max())
    o = G.triangulate(g[tag + "_tri_P1"], g[tag + "_tri_P2"], p1[sel], p2[sel])
    assert np.allclose(a, o[:3] / o[3], rtol=1e-4, atol=1e-4 * np.abs(b).max())


@pytest.mark.parametrize("name,tag,K", CASES)
def test_ransac_estimators_agree_with_reference(name, tag, K):
    from orbslam2 import utils
    g = load_golden(name)
    p1, p2 = g[tag + "_p1"], g[tag + "_p2"]
    F, mask = utils.find_fundamental_ransac(p1, p2, 3.0, 0.99)
    ref = g[tag + "_Fmask"]
    assert F.shape == (3, 3) and abs(F[2, 2] - 1.0) < 1e-12 and mask.shape == (len(p1), 1)
    assert _jaccard(mask, ref) >= 0.93, _jaccard(mask, ref)
    assert int(mask.sum()) >= 0.97 * int(ref.sum())
    assert abs(np.linalg.det(F)) < 1e-6 * np.abs(F).max() ** 3 + 1e-12            # rank 2
    assert np.array_equal(utils.score_fundamental(F, p1, p2, 3.0), mask.ravel())   # mask belongs to the returned F
    F2, mask2 = utils.find_fundamental_ransac(p1, p2, 3.0, 0.99)                   # deterministic for a seed
    assert np.array_equal(F, F2) and np.array_equal(mask, mask2)
    for thr in (1.0, 3.0):
        E, emask = utils.calculate_essential_matrix(p1, p2, K, prob=0.999, threshold=thr)
        ref = g["%s_Emask_%d" % (tag, int(thr))]
        jac = _jaccard(emask, ref)
        assert jac >= (0.97 if thr == 3.0 else 0.80), (thr, jac, int(emask.sum()), int(ref.sum()))
        assert int(emask.sum()) >= 0.97 * int(ref.sum())
        s = np.linalg.svd(E, compute_uv=False)
        assert abs(s[0] - s[1]) < 1e-6 and s[2] < 1e-6 and abs(s[0] - np.sqrt(0.5)) < 1e-6
        assert np.array_equal(utils.score_essential(E, p1, p2, K, thr), emask.ravel())


def test_ransac_small_and_degenerate_inputs():
    from orbslam2 import _native, utils
    rng = np.random.default_rng(1)
    p1 = rng.uniform(50, 400, (8, 2)).astype(np.float32)
    p2 = p1 + rng.normal(0, 1, (8, 2)).astype(np.float32)
    F, mask = utils.find_fundamental_ransac(p1, p2, 3.0, 0.99)
    assert mask.shape == (8, 1) and (F is None or F.shape == (3, 3))
    with pytest.raises(_native.VslamError):
        utils.find_fundamental_ransac(p1[:6], p2[:6])
    with pytest.raises(_native.VslamError):
        utils.calculate_essential_matrix(p1[:4], p2[:4], np.eye(3))
    X = utils.triangulate_points(np.zeros((0, 2), np.float32), np.zeros((0, 2), np.float32), np.eye(3, 4), np.eye(3, 4))
    assert X.shape == (4, 0)
