"""CPU tests: the C-ABI library loads and exports every symbol include/vslam_b200.h declares
(no compute calls without a GPU), and the host-side logic of the drop-in classes."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared_functions():
    text = open(os.path.join(ROOT, "include", "vslam_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vslam_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from orbslam2 import _native
    assert os.path.exists(_native.LIB_PATH), "run `python visual-slam_b200/build.py` (or __graft_entry__.build())"
    lib = ctypes.CDLL(_native.LIB_PATH)
    names = _declared_functions()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    lib.vslam_abi_version.restype = ctypes.c_int
    assert lib.vslam_abi_version() >= 1000
    lib.vslam_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.vslam_last_error(), bytes)


def test_invalid_arguments_are_status_codes_not_crashes():
    from orbslam2 import _native
    L = _native.lib()
    assert L.vslam_orb_create(None, None) == -1
    assert b"null" in L.vslam_last_error()
    assert L.vslam_orb_destroy(None) == 0
    assert L.vslam_orb_kp_capacity(None) == -1
    assert L.vslam_bf_knn2_hamming256_host(None, -1, None, 0, None, None) == -1


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "visual-slam_b200", "orbslam2")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert "oracle" not in src.replace("no CPU fallback (the CPU restatement under\noracle/", ""), f


def test_no_gpu_means_loud_failure():
    from orbslam2 import _native
    if _native.lib().vslam_device_count() > 0:
        pytest.skip("a GPU is present")
    from orbslam2.extractor import ORBExtractor
    from orbslam2.matcher import DescriptorMatcher
    with pytest.raises(_native.VslamError):
        ORBExtractor().detect_and_compute(np.zeros((480, 640), np.uint8))
    with pytest.raises(_native.VslamError):
        DescriptorMatcher().match(np.zeros((4, 32), np.uint8), np.zeros((4, 32), np.uint8))


def test_matcher_host_logic():
    cv2 = pytest.importorskip("cv2")
    from orbslam2.matcher import DescriptorMatcher
    m = DescriptorMatcher()
    assert m.ratio_threshold == 0.75
    assert m.match(None, np.zeros((3, 32), np.uint8)) == []
    assert m.match(np.zeros((0, 32), np.uint8), np.zeros((3, 32), np.uint8)) == []
    assert m.match_with_mask(None, None, None) == []
    with pytest.raises(ValueError):
        DescriptorMatcher("nope")
    with pytest.raises(NotImplementedError):
        DescriptorMatcher("flann")
    ms = [cv2.DMatch(0, 0, 0, 30.0), cv2.DMatch(1, 1, 0, 10.0), cv2.DMatch(2, 2, 0, 70.0), cv2.DMatch(3, 3, 0, 10.0)]
    out = m.filter_matches_by_distance(ms)                     # median 20 -> keep < 40, stable order
    assert [(x.queryIdx, x.distance) for x in out] == [(1, 10.0), (3, 10.0), (0, 30.0)]
    assert m.filter_matches_by_distance([]) == []
    kp1 = [cv2.KeyPoint(10.0 * i, 0.0, 31) for i in range(4)]
    kp2 = [cv2.KeyPoint(10.0 * i + (0, 5, 50, 200)[i], 0.0, 31) for i in range(4)]
    out = m.filter_matches_by_geometric_distance(kp1, kp2, ms, 0.02, (480, 640))   # max_dist 11.2
    assert [x.queryIdx for x in out] == [0, 1]
    few, mask = m.filter_matches_by_fundamental(kp1, kp2, ms)
    assert few is ms and mask.dtype == bool and mask.all() and len(mask) == 4


def test_utils_host_logic():
    from orbslam2 import utils
    K = np.array([[100.0, 0, 50], [0, 100, 40], [0, 0, 1]])
    R = np.eye(3)
    for t in (np.array([1.0, 2, 3]), np.array([[1.0, 2, 3]]), np.array([[1.0], [2], [3]])):
        P = utils.compute_projection_matrix(R, t, K)
        assert P.shape == (3, 4) and np.allclose(P[:, 3], K @ np.array([1.0, 2, 3]))
    X = np.array([[2.0, 4], [4, 8], [6, 12], [2, 4]])
    assert np.allclose(utils.convert_to_3d_points(X), [[1, 2, 3], [1, 2, 3]])
    assert np.allclose(utils.convert_to_3d_points(X.T), [[1, 2, 3], [1, 2, 3]])
    assert utils.RANSAC == 8


def test_initializer_guards():
    from orbslam2.initializer import MapInitializer
    init = MapInitializer(np.eye(3), min_matches=10)
    assert init.initialize([], None, None, None) == (False, None, None, None, None)
    assert init.initialization_done is False and init.first_frame_keypoints is None

    class FewMatches:
        def match(self, a, b):
            return [1, 2, 3]
    init.set_first_frame([1], np.zeros((1, 32), np.uint8), np.zeros((4, 4), np.uint8))
    ok, R, t, pts, ms = init.initialize([1], np.zeros((1, 32), np.uint8), FewMatches(), None)
    assert ok is False and ms == [1, 2, 3]


def test_frontend_gates_equal_the_reference_loops():
    """frontend.gate_* (array form of matcher.py:109-169) against the reference's Python loops, including
    displacements that sit exactly on the threshold and heavy distance ties."""
    import math
    from orbslam2 import frontend
    rng = np.random.default_rng(3)
    n = 400
    pts1 = rng.uniform(0, 640, (n, 2)).astype(np.float32)
    pts2 = (pts1 + rng.normal(0, 8, (n, 2))).astype(np.float32)
    pts2[:40] = pts1[:40] + np.float32([11.2, 0.0])           # |d| == max_dist for (480, 640) at 2 %  (11.2)
    pts2[40:60] = pts1[40:60] + np.float32([6.72, 8.96])      # 3-4-5 triangle scaled to 11.2
    q = rng.permutation(n)[:300].astype(np.int32)
    q.sort()
    t = rng.integers(0, n, 300).astype(np.int32)
    t[:60] = q[:60]
    dist = rng.integers(5, 60, 300).astype(np.float32)
    max_dist = ((640 + 480) / 2.0) * 0.02
    ref_keep = [math.hypot(float(pts2[b][0]) - float(pts1[a][0]), float(pts2[b][1]) - float(pts1[a][1])) <= max_dist
                for a, b in zip(q, t)]
    keep = frontend.gate_geometric_distance(pts1, pts2, q, t, 0.02, (480, 640))
    assert keep.tolist() == ref_keep and 0 < sum(ref_keep) < 300
    for d in (dist, dist[:1], dist[:2], np.full(7, 9, np.float32)):
        order, n_keep = frontend.gate_descriptor_distance(d)
        ref_sorted = sorted(range(len(d)), key=lambda i: d[i])
        thr = np.median([float(x) for x in d]) * 2.0
        ref = [i for i in ref_sorted if d[i] < thr]
        assert order[:n_keep].tolist() == ref
    order, n_keep = frontend.gate_descriptor_distance(dist, 20.0)
    assert order[:n_keep].tolist() == [i for i in sorted(range(300), key=lambda i: dist[i]) if dist[i] < 20.0]
    assert frontend.gate_descriptor_distance(np.zeros(0, np.float32)) [1] == 0
