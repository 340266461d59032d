"""ctypes front end of the CPU oracle (oracle/*.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs -- never by the product package
(visual-slam_b200/), which must fail loudly when its CUDA library is missing.

Each function cites the reference call it restates; the arithmetic notes are in
the C sources.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".inc"))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_harris_response.restype = C.c_float
        _LIB.orc_ic_angle.restype = C.c_float
        _LIB.orc_fast_atan2.restype = C.c_float
        _LIB.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
    return _LIB


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


# ---------------------------------------------------------------- extraction
def bgr2gray(bgr):
    """extractor.py:61-62 cv2.cvtColor(image, COLOR_BGR2GRAY)."""
    bgr = _u8(bgr)
    h, w, _ = bgr.shape
    out = np.empty((h, w), np.uint8)
    lib().orc_bgr2gray(bgr.ctypes.data_as(C.c_void_p), h, w, out.ctypes.data_as(C.c_void_p))
    return out


def level_geometry(w, h, scale_factor=1.2, nlevels=8):
    lw = np.zeros(nlevels, np.int32)
    lh = np.zeros(nlevels, np.int32)
    ls = np.zeros(nlevels, np.float32)
    lib().orc_level_geometry(int(w), int(h), C.c_float(np.float32(scale_factor)), nlevels,
                             lw.ctypes.data_as(C.c_void_p), lh.ctypes.data_as(C.c_void_p),
                             ls.ctypes.data_as(C.c_void_p))
    return lw, lh, ls


def level_quotas(nfeatures, scale_factor=1.2, nlevels=8):
    q = np.zeros(nlevels, np.int32)
    lib().orc_level_quotas(int(nfeatures), C.c_float(np.float32(scale_factor)), nlevels,
                           q.ctypes.data_as(C.c_void_p))
    return q


def resize_linear_exact(src, dw, dh):
    """cv2.resize(src, (dw, dh), interpolation=INTER_LINEAR_EXACT) for uint8."""
    src = _u8(src)
    sh, sw = src.shape
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_exact(src.ctypes.data_as(C.c_void_p), sw, sh, sw,
                                  dst.ctypes.data_as(C.c_void_p), dw, dh, dw)
    return dst


def linear_exact_table(src_n, dst_n):
    i0 = np.zeros(dst_n, np.int32)
    aq = np.zeros(dst_n, np.int32)
    lib().orc_linear_exact_table(int(src_n), int(dst_n), i0.ctypes.data_as(C.c_void_p),
                                 aq.ctypes.data_as(C.c_void_p))
    return i0, aq


def fast_detect(img, thr=7, nms=True):
    """cv2.FastFeatureDetector_create(thr, nms).detect -> (x, y, score) in raster order."""
    img = _u8(img)
    h, w = img.shape
    cap = w * h // 2 + 16
    xs = np.zeros(cap, np.int32)
    ys = np.zeros(cap, np.int32)
    ss = np.zeros(cap, np.int32)
    n = lib().orc_fast_detect(img.ctypes.data_as(C.c_void_p), w, h, w, int(thr), int(bool(nms)),
                              xs.ctypes.data_as(C.c_void_p), ys.ctypes.data_as(C.c_void_p),
                              ss.ctypes.data_as(C.c_void_p), cap)
    return xs[:n].copy(), ys[:n].copy(), ss[:n].copy()


def fast_score_map(img, thr=7):
    img = _u8(img)
    h, w = img.shape
    out = np.zeros((h, w), np.uint8)
    lib().orc_fast_score_map(img.ctypes.data_as(C.c_void_p), w, h, w, int(thr), out.ctypes.data_as(C.c_void_p))
    return out


def gaussian7(img):
    """The 7x7 sigma=2 blur cv2.ORB applies to each level before rBRIEF."""
    img = _u8(img)
    h, w = img.shape
    out = np.empty_like(img)
    lib().orc_gaussian7(img.ctypes.data_as(C.c_void_p), w, h, w, out.ctypes.data_as(C.c_void_p), w)
    return out


def harris_response(img, x, y):
    img = _u8(img)
    return float(lib().orc_harris_response(img.ctypes.data_as(C.c_void_p), img.shape[1], int(x), int(y)))


def ic_angle(img, x, y):
    img = _u8(img)
    return float(lib().orc_ic_angle(img.ctypes.data_as(C.c_void_p), img.shape[1], int(x), int(y)))


def pyramid_level(img, level, scale_factor=1.2, nlevels=8, blurred=False):
    img = _u8(img)
    h, w = img.shape
    out = np.empty(h * w, np.uint8)
    ow, oh = C.c_int(), C.c_int()
    n = lib().orc_pyramid_level(img.ctypes.data_as(C.c_void_p), w, h, C.c_float(np.float32(scale_factor)),
                                nlevels, level, int(blurred), out.ctypes.data_as(C.c_void_p),
                                C.byref(ow), C.byref(oh))
    assert n >= 0
    return out[:n].reshape(oh.value, ow.value).copy()


def level_fast_candidates(img, level, scale_factor=1.2, nlevels=8, fast_thr=7):
    img = _u8(img)
    h, w = img.shape
    cap = w * h // 4 + 16
    xs = np.zeros(cap, np.int32)
    ys = np.zeros(cap, np.int32)
    ss = np.zeros(cap, np.int32)
    n = lib().orc_level_fast_candidates(img.ctypes.data_as(C.c_void_p), h, w,
                                        C.c_float(np.float32(scale_factor)), nlevels, level, int(fast_thr),
                                        xs.ctypes.data_as(C.c_void_p), ys.ctypes.data_as(C.c_void_p),
                                        ss.ctypes.data_as(C.c_void_p), cap)
    return xs[:n].copy(), ys[:n].copy(), ss[:n].copy()


def orb_detect_and_compute(img, nfeatures=2000, scale_factor=1.2, nlevels=8, fast_thr=7, cap=None):
    """extractor.py:50-67 ORBExtractor.detect_and_compute on a gray image.

    Returns a dict of arrays: x, y, size, angle, response (float32), octave, lx, ly
    (int32), desc uint8[n,32], level_counts int32[nlevels].  Order: level-major,
    (y, x) ascending inside a level (cv2's own in-level order is unspecified).
    """
    img = _u8(img)
    h, w = img.shape
    cap = int(cap or max(4 * nfeatures + 64, 256))
    f = lambda: np.zeros(cap, np.float32)
    i = lambda: np.zeros(cap, np.int32)
    x, y, size, ang, resp = f(), f(), f(), f(), f()
    octv, lx, ly = i(), i(), i()
    desc = np.zeros((cap, 32), np.uint8)
    lc = np.zeros(nlevels, np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    n = lib().orc_orb_detect_and_compute(p(img), h, w, int(nfeatures), C.c_float(np.float32(scale_factor)),
                                         int(nlevels), int(fast_thr), cap, p(x), p(y), p(size), p(ang),
                                         p(resp), p(octv), p(lx), p(ly), p(desc), p(lc))
    if n < 0:
        raise ValueError("oracle: bad parameters")
    if n > cap:
        return orb_detect_and_compute(img, nfeatures, scale_factor, nlevels, fast_thr, cap=n + 64)
    return dict(x=x[:n].copy(), y=y[:n].copy(), size=size[:n].copy(), angle=ang[:n].copy(),
                response=resp[:n].copy(), octave=octv[:n].copy(), lx=lx[:n].copy(), ly=ly[:n].copy(),
                desc=desc[:n].copy(), level_counts=lc)


def orb_compute(img, kx, ky, kangle, koct, scale_factor=1.2):
    """extractor.py:68-83 ORBExtractor.compute -> (desc uint8[m,32], keep_index int32[m])."""
    img = _u8(img)
    h, w = img.shape
    kx = np.ascontiguousarray(kx, np.float32)
    ky = np.ascontiguousarray(ky, np.float32)
    ka = np.ascontiguousarray(kangle, np.float32)
    ko = np.ascontiguousarray(koct, np.int32)
    n = len(kx)
    desc = np.zeros((max(n, 1), 32), np.uint8)
    keep = np.zeros(max(n, 1), np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    m = lib().orc_orb_compute(p(img), h, w, C.c_float(np.float32(scale_factor)), p(kx), p(ky), p(ka), p(ko),
                              n, p(desc), p(keep))
    if m < 0:
        raise ValueError("oracle: bad parameters")
    return desc[:m].copy(), keep[:m].copy()


# ------------------------------------------------------------------ matching
def bf_knn2(q, t):
    """matcher.py:70 knnMatch(d1, d2, k=2) -> idx int32[nq,2], dist int32[nq,2] (-1 = missing)."""
    q = _u8(q).reshape(-1, 32)
    t = _u8(t).reshape(-1, 32)
    idx = np.full((len(q), 2), -1, np.int32)
    dist = np.full((len(q), 2), -1, np.int32)
    if len(q):
        lib().orc_bf_knn2(q.ctypes.data_as(C.c_void_p), len(q), t.ctypes.data_as(C.c_void_p), len(t),
                          idx.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p))
    return idx, dist


def bf_match_masked(q, t, mask=None):
    """matcher.py:102 BFMatcher.match(d1, d2, mask=mask) (k=1)."""
    q = _u8(q).reshape(-1, 32)
    t = _u8(t).reshape(-1, 32)
    idx = np.full(len(q), -1, np.int32)
    dist = np.full(len(q), -1, np.int32)
    m = None if mask is None else _u8(mask)
    if len(q):
        lib().orc_bf_match_masked(q.ctypes.data_as(C.c_void_p), len(q), t.ctypes.data_as(C.c_void_p), len(t),
                                  None if m is None else m.ctypes.data_as(C.c_void_p),
                                  idx.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p))
    return idx, dist


def ratio_filter(idx, dist, ratio=0.75, ratio_test=True):
    """matcher.py:73-81 ratio loop -> bool keep[nq]."""
    idx = np.ascontiguousarray(idx, np.int32)
    dist = np.ascontiguousarray(dist, np.int32)
    keep = np.zeros(len(idx), np.uint8)
    if len(idx):
        lib().orc_ratio_filter(idx.ctypes.data_as(C.c_void_p), dist.ctypes.data_as(C.c_void_p), len(idx),
                               C.c_double(ratio), int(bool(ratio_test)), keep.ctypes.data_as(C.c_void_p))
    return keep.astype(bool)


def match(d1, d2, ratio=0.75, ratio_test=True):
    """matcher.py:44-83 DescriptorMatcher.match -> list of (queryIdx, trainIdx, distance)."""
    if d1 is None or d2 is None or len(d1) == 0 or len(d2) == 0:
        return []
    idx, dist = bf_knn2(np.uint8(d1), np.uint8(d2))
    keep = ratio_filter(idx, dist, ratio, ratio_test)
    return [(int(i), int(idx[i, 0]), float(dist[i, 0])) for i in np.nonzero(keep)[0]]


# ------------------------------------------------------------- distribute_keypoints
def min_eigen_map(img):
    """cv2.cornerMinEigenVal(img, 3, ksize=3) (inside goodFeaturesToTrack, extractor.py:123-129)."""
    img = _u8(img)
    h, w = img.shape
    out = np.zeros((h, w), np.float32)
    lib().orc_min_eigen_map(img.ctypes.data_as(C.c_void_p), w, h, out.ctypes.data_as(C.c_void_p))
    return out


def gftt_grid(img, n_features=2000, rows=8, cols=8, quality=0.01, min_distance=10.0):
    """extractor.py:85-136: corner positions of distribute_keypoints -> (xy float32 [n,2], cell_count)."""
    img = _u8(img)
    h, w = img.shape
    per = n_features // (rows * cols)
    xy = np.zeros((max(rows * cols * per, 1), 2), np.float32)
    cc = np.zeros(rows * cols, np.int32)
    n = lib().orc_gftt_grid(img.ctypes.data_as(C.c_void_p), h, w, rows, cols, per, C.c_double(quality),
                            C.c_double(min_distance), xy.ctypes.data_as(C.c_void_p), cc.ctypes.data_as(C.c_void_p))
    return xy[:n].copy(), cc
