"""Drop-in geometry helpers (reference: src/orbslam2/utils.py:55-70,120-160) on the B200 kernels.

Same names, signatures and return shapes/dtypes as the reference functions, which are thin
wrappers over cv2.findEssentialMat / recoverPose / triangulatePoints.  The config / IO
helpers of the reference's utils.py (load_config, PLY writer, ...) are host-side glue outside
the hot path and are not re-implemented here.
"""
import ctypes as C

import numpy as np

from . import _native

RANSAC = 8          # value of cv2.RANSAC, so that callers can pass method=cv2.RANSAC unchanged
# seed 0 = OpenCV's own sample stream (cv::RNG(-1) on every call; the reference cannot change it: cv2's estimators
# ignore cv2.setRNGSeed).  Any other value gives a different deterministic stream.
DEFAULT_SEED = 0


def _pts(p):
    return np.ascontiguousarray(np.asarray(p, np.float32).reshape(-1, 2))


def find_fundamental_ransac(points1, points2, threshold=3.0, confidence=0.99, max_iters=1000, seed=DEFAULT_SEED):
    """cv2.findFundamentalMat(p1, p2, FM_RANSAC, threshold, confidence) -> (F float64 [3,3], mask uint8 [n,1])."""
    p1, p2 = _pts(points1), _pts(points2)
    n = len(p1)
    _native.require_gpu()
    F = np.zeros(9, np.float64)
    mask = np.zeros(n, np.uint8)
    cnt = C.c_int()
    _native.check(_native.lib().vslam_ransac_fundamental_host(_native.ptr(p1), _native.ptr(p2), n, float(threshold),
                                                              float(confidence), int(max_iters), C.c_uint64(seed),
                                                              _native.ptr(F), _native.ptr(mask), C.byref(cnt)))
    if not F.any():                  # no model (cv2 returns None); the LMedS path (n < 15) still reports its mask
        return None, mask.reshape(-1, 1)
    return F.reshape(3, 3), mask.reshape(-1, 1)


def score_fundamental(F, points1, points2, threshold=3.0):
    """Inlier mask of a given F under OpenCV's FM_RANSAC error (max of the two squared epipolar distances)."""
    p1, p2 = _pts(points1), _pts(points2)
    _native.require_gpu()
    Fm = np.ascontiguousarray(F, np.float64).reshape(9)
    mask = np.zeros(len(p1), np.uint8)
    cnt = C.c_int()
    _native.check(_native.lib().vslam_score_fundamental_host(_native.ptr(Fm), _native.ptr(p1), _native.ptr(p2), len(p1),
                                                             float(threshold), _native.ptr(mask), C.byref(cnt)))
    return mask


def score_essential(E, points1, points2, camera_matrix, threshold=1.0):
    """Inlier mask of a given E under OpenCV's Sampson error in K-normalised coordinates."""
    p1, p2 = _pts(points1), _pts(points2)
    _native.require_gpu()
    Em = np.ascontiguousarray(E, np.float64).reshape(9)
    K = np.ascontiguousarray(camera_matrix, np.float64).reshape(9)
    mask = np.zeros(len(p1), np.uint8)
    cnt = C.c_int()
    _native.check(_native.lib().vslam_score_essential_host(_native.ptr(Em), _native.ptr(K), _native.ptr(p1), _native.ptr(p2),
                                                           len(p1), float(threshold), _native.ptr(mask), C.byref(cnt)))
    return mask


def undistort_image(image, camera_matrix, distortion):
    """utils.py:40-52 `cv2.undistort(image, camera_matrix, distortion)` on the device: uint8 gray or BGR image
    -> undistorted image of the same shape (bit-exact with OpenCV's fixed-point remap)."""
    image = np.asarray(image)
    if image.dtype != np.uint8 or image.ndim not in (2, 3) or (image.ndim == 3 and image.shape[2] not in (1, 3)):
        raise ValueError("undistort_image: uint8 image with 1 or 3 channels expected")
    src = np.ascontiguousarray(image)
    h, w = src.shape[:2]
    ch = 1 if src.ndim == 2 else src.shape[2]
    K = np.ascontiguousarray(camera_matrix, np.float64).reshape(3, 3)
    dist = np.zeros(0) if distortion is None else np.ascontiguousarray(distortion, np.float64).ravel()
    out = np.empty_like(src)
    _native.require_gpu()
    _native.check(_native.lib().vslam_undistort_host(_native.ptr(src), h, w, ch, _native.ptr(K),
                                                     _native.ptr(dist) if len(dist) else None, len(dist),
                                                     _native.ptr(out)))
    return out


def triangulate_points(points1, points2, P1, P2):
    """utils.py:56-60 -> float32 [4, N] homogeneous points (as cv2.triangulatePoints returns them)."""
    p1, p2 = _pts(points1), _pts(points2)
    n = len(p1)
    _native.require_gpu()
    A = np.ascontiguousarray(P1, np.float64).reshape(12)
    B = np.ascontiguousarray(P2, np.float64).reshape(12)
    X = np.zeros((4, n), np.float32)
    _native.check(_native.lib().vslam_triangulate_host(_native.ptr(A), _native.ptr(B), _native.ptr(p1), _native.ptr(p2), n,
                                                       _native.ptr(X)))
    return X


def convert_to_3d_points(points_4d):
    """utils.py:62-70: homogeneous (x, y, z, w) -> Euclidean [N, 3]."""
    if points_4d.shape[0] != 4:
        points_4d = points_4d.T
    return (points_4d[:3] / points_4d[3]).T


def calculate_essential_matrix(points1, points2, camera_matrix, method=RANSAC, prob=0.999, threshold=1.0,
                               max_iters=1000, seed=DEFAULT_SEED):
    """utils.py:120-126 -> (E float64 [3,3], mask uint8 [n,1])."""
    if method != RANSAC:
        raise NotImplementedError("only method=cv2.RANSAC (8) is on the B200 hot path")
    p1, p2 = _pts(points1), _pts(points2)
    n = len(p1)
    _native.require_gpu()
    K = np.ascontiguousarray(camera_matrix, np.float64).reshape(9)
    E = np.zeros(9, np.float64)
    mask = np.zeros(n, np.uint8)
    cnt = C.c_int()
    _native.check(_native.lib().vslam_ransac_essential_host(_native.ptr(p1), _native.ptr(p2), n, _native.ptr(K), float(prob),
                                                            float(threshold), int(max_iters), C.c_uint64(seed),
                                                            _native.ptr(E), _native.ptr(mask), C.byref(cnt)))
    if cnt.value == 0:
        return None, mask.reshape(-1, 1)
    return E.reshape(3, 3), mask.reshape(-1, 1)


def recover_pose(E, points1, points2, camera_matrix, mask=None):
    """utils.py:129-134 -> (n_inliers, R float64 [3,3], t float64 [3,1], pose_mask uint8 [n,1])."""
    p1, p2 = _pts(points1), _pts(points2)
    n = len(p1)
    _native.require_gpu()
    Em = np.ascontiguousarray(E, np.float64).reshape(9)
    K = np.ascontiguousarray(camera_matrix, np.float64).reshape(9)
    m = np.ones(n, np.uint8) if mask is None else np.ascontiguousarray(np.asarray(mask).reshape(-1) != 0, np.uint8)
    R = np.zeros(9, np.float64)
    t = np.zeros(3, np.float64)
    good = C.c_int()
    _native.check(_native.lib().vslam_recover_pose_host(_native.ptr(Em), _native.ptr(p1), _native.ptr(p2), n, _native.ptr(K),
                                                        _native.ptr(m), _native.ptr(R), _native.ptr(t), C.byref(good)))
    return good.value, R.reshape(3, 3), t.reshape(3, 1), m.reshape(-1, 1)


def compute_projection_matrix(R, t, camera_matrix):
    """utils.py:137-160: P = K [R | t] with t accepted as (3,), (1,3) or (3,1)."""
    t = np.array(t)
    if t.ndim == 1:
        t = t.reshape(3, 1)
    elif t.ndim == 2 and t.shape[0] == 1 and t.shape[1] == 3:
        t = t.reshape(3, 1)
    return camera_matrix @ np.hstack((R, t))
