"""GPU parity of vslam_cull_map_points (LocalMapper._cull_map_points, local_mapper.py:205-245) through the C ABI:
the kept map points equal the reference's own on the golden maps (tests/golden/cull.npz), and the decisions /
errors equal the oracle's on a large seeded map.

Tolerance (float64 arithmetic whose summation order differs from numpy's BLAS-backed `P @ x`): errors agree to
1e-9 px absolute; a decision may differ only where the error is within 1e-9 px of the 5 px gate (none of the
golden or seeded cases is)."""
import os
import types

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cull.npz")
TOL = 1e-9


def _case(g, c):
    p = "c%d_" % c
    d = {k[len(p):]: g[k] for k in g.files if k.startswith(p)}
    d["obs_xy"] = d["kp_xy"][d["kp_off"][d["obs_kf"]] + d["obs_kp"]] if len(d["obs_kf"]) else np.zeros((0, 2), np.float32)
    return d


def test_cull_equals_reference_golden():
    from orbslam2 import frontend
    from oracle import cull_oracle
    g = np.load(GOLD)
    for c in range(int(g["n"])):
        d = _case(g, c)
        P = cull_oracle.projection_matrices(d["K"], d["poses"])
        valid, err = frontend.reprojection_cull(d["points"], d["offs"], d["obs_kf"], d["obs_xy"], P, int(d["min_obs"]))
        assert np.array_equal(np.flatnonzero(valid).astype(np.int32), d["kept"]), c
        ov, oe = cull_oracle.cull(d["points"], d["offs"], d["obs_kf"], d["obs_xy"], P, int(d["min_obs"]))
        assert np.array_equal(valid, ov)
        fin = np.isfinite(oe)
        assert np.array_equal(np.isfinite(err), fin) and np.allclose(err[fin], oe[fin], rtol=0, atol=TOL)


def test_local_mapper_method_on_reference_structures():
    """A class laid out like the reference's LocalMapper (keyframe / map-point dicts, cv2.KeyPoint objects) gets
    its _cull_map_points replaced by install_local_mapper_adaptor and ends with the reference's map."""
    import cv2
    from orbslam2 import frontend
    g = np.load(GOLD)
    mod = types.ModuleType("fake_local_mapper")
    mod.cv2 = cv2

    class LocalMapper:
        def _cull_map_points(self):
            raise AssertionError("the adaptor must replace this method")
    mod.LocalMapper = LocalMapper
    proxy = frontend.install_local_mapper_adaptor(mod)
    try:
        for c in range(int(g["n"])):
            d = _case(g, c)
            lm = mod.LocalMapper()
            lm.camera_matrix, lm.min_observations = d["K"], int(d["min_obs"])
            lm.keyframes = []
            for k, T in enumerate(d["poses"]):
                xy = d["kp_xy"][d["kp_off"][k]:d["kp_off"][k + 1]]
                lm.keyframes.append(dict(id=k, keypoints=[cv2.KeyPoint(float(x), float(y), 31.0) for x, y in xy],
                                         pose=T, map_points=[]))
            lm.map_points = []
            for i in range(len(d["points"])):
                obs = {int(d["obs_kf"][j]): int(d["obs_kp"][j]) for j in range(d["offs"][i], d["offs"][i + 1])}
                lm.map_points.append(dict(id=i, position=d["points"][i], observed_keyframes=obs))
            lm._cull_map_points()
            assert [mp["id"] for mp in lm.map_points] == d["kept"].tolist(), c
            for k, kf in enumerate(lm.keyframes):
                assert kf["map_points"] == d["kf_lists"][d["kf_lists_off"][k]:d["kf_lists_off"][k + 1]].tolist()
    finally:
        mod.LocalMapper._cull_map_points = proxy._cull
    assert frontend.cull_map_points([], [], np.eye(3)) == []


def test_cull_large_map_equals_oracle_and_guards():
    from orbslam2 import _native, frontend
    from oracle import cull_oracle
    rng = np.random.default_rng(5)
    n_kf, n = 40, 20000
    poses = np.tile(np.eye(4), (n_kf, 1, 1))
    poses[:, :3, 3] = rng.normal(0, 0.5, (n_kf, 3))
    K = np.array([[700.0, 0, 640], [0, 705.0, 360], [0, 0, 1]])
    P = cull_oracle.projection_matrices(K, poses)
    pts = np.c_[rng.uniform(-5, 5, n), rng.uniform(-4, 4, n), rng.uniform(3, 20, n)]
    cnt = rng.integers(0, 9, n)
    off = np.r_[0, np.cumsum(cnt)].astype(np.int32)
    kf = rng.integers(0, n_kf, off[-1]).astype(np.int32)
    rep = np.repeat(np.arange(n), cnt)
    h = np.einsum("mij,mj->mi", P[kf], np.c_[pts[rep], np.ones(len(rep))])
    xy = (h[:, :2] / h[:, 2:3] + rng.normal(0, 2.2, (len(rep), 2))).astype(np.float32)
    valid, err = frontend.reprojection_cull(pts, off, kf, xy, P, 2, 5.0)
    ov, oe = cull_oracle.cull(pts, off, kf, xy, P, 2, 5.0)
    assert np.abs(oe - 5.0).min() > TOL            # no borderline decision in this map
    assert np.array_equal(valid, ov) and 0.2 < valid.mean() < 0.8
    assert np.allclose(err, oe, rtol=0, atol=TOL)
    # other gates and counts
    v3, _ = frontend.reprojection_cull(pts, off, kf, xy, P, 4, 3.0)
    o3, _ = cull_oracle.cull(pts, off, kf, xy, P, 4, 3.0)
    assert np.array_equal(v3, o3)
    # an observation of a keyframe that does not exist: the reference raises IndexError, the library refuses
    bad = kf.copy()
    bad[3] = n_kf
    with pytest.raises(_native.VslamError):
        frontend.reprojection_cull(pts, off, bad, xy, P, 0, 1e30)
    with pytest.raises(ValueError):
        frontend.reprojection_cull(pts, off[:-1], kf, xy, P)
    e, _ = frontend.reprojection_cull(np.zeros((0, 3)), [0], [], np.zeros((0, 2)), P)
    assert e.shape == (0,)
