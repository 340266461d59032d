"""ctypes binding of libvslam_b200.so (C ABI in include/vslam_b200.h).

The CUDA library is the only compute path of this package: if it is missing or no
B200 is visible, calls raise -- there is no CPU fallback (the CPU restatement under
oracle/ is test infrastructure and is never imported from here).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# VSLAM_LIB_PATH: a differently built libvslam_b200 (kernel A/B runs, tools/build_variant.py); same ABI, same loud failure
LIB_PATH = os.environ.get("VSLAM_LIB_PATH") or os.path.join(_HERE, "libvslam_b200.so")
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4")])


class OrbParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("n_features", C.c_int32),
                ("n_levels", C.c_int32), ("fast_threshold", C.c_int32), ("scale_factor", C.c_float),
                ("max_batch", C.c_int32), ("kp_capacity", C.c_int32)]


class VslamError(RuntimeError):
    pass


def lib():
    """Load the CUDA library (fails loudly when it has not been built)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise VslamError(
                "libvslam_b200.so is not built (%s); run `python visual-slam_b200/build.py` -- "
                "this package has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.vslam_last_error.restype = C.c_char_p
        vp, i32, f64, i64 = C.c_void_p, C.c_int, C.c_double, C.c_long
        L.vslam_orb_create.argtypes = [C.POINTER(OrbParams), C.POINTER(vp)]
        L.vslam_orb_destroy.argtypes = [vp]
        L.vslam_orb_kp_capacity.argtypes = [vp]
        L.vslam_orb_set_distribution.argtypes = [vp, i32, i32]
        L.vslam_orb_level_info.argtypes = [vp, i32] + [vp] * 5
        L.vslam_orb_detect_and_compute.argtypes = [vp, vp, i32, i32, vp, vp, vp, vp]
        L.vslam_orb_detect_and_compute_host.argtypes = [vp, vp, i32, vp, vp, vp]
        L.vslam_orb_detect_and_compute_host_bgr.argtypes = [vp, vp, i32, vp, vp, vp]
        L.vslam_bgr2gray.argtypes = [vp, C.c_size_t, vp, vp]
        L.vslam_undistort.argtypes = [vp, i32, i32, i32, i32, i32, vp, vp, i32, vp, i32, vp]
        L.vslam_undistort_host.argtypes = [vp, i32, i32, i32, vp, vp, i32, vp]
        L.vslam_orb_last_status.argtypes = [vp]
        L.vslam_orb_debug_level.argtypes = [vp, i32, i32, i32, vp]
        L.vslam_orb_debug_candidates.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp]
        L.vslam_bf_knn2_hamming256.argtypes = [vp, i32, i64, vp, i32, i64, vp, vp, i32, vp, vp, vp]
        L.vslam_bf_knn2_hamming256_host.argtypes = [vp, i32, vp, i32, vp, vp]
        L.vslam_bf_match_masked_host.argtypes = [vp, i32, vp, i32, vp, vp, vp]
        L.vslam_ratio_filter.argtypes = [vp, vp, i32, vp, i32, f64, i32, vp, vp, vp]
        L.vslam_ransac_fundamental_batched_workspace.argtypes = [i32, i32]
        L.vslam_ransac_fundamental_batched_workspace.restype = C.c_size_t
        L.vslam_ransac_fundamental_batched.argtypes = [vp, vp, i32, vp, vp, i32, i32, f64, f64, i32, C.c_uint64, vp, vp, vp, vp, vp]
        L.vslam_track_batched_workspace.argtypes = [i32, i32]
        L.vslam_track_batched_workspace.restype = C.c_size_t
        L.vslam_track_batched.argtypes = [vp, vp, i32, vp, vp, i32, i32, vp, i32, i32, f64, f64, f64, i32, C.c_uint64,
                                          vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L.vslam_track_clip_host.argtypes = [vp, vp, i32, f64, vp, f64, f64, f64, i32, C.c_uint64, vp, vp, vp]
        L.vslam_cull_map_points.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, f64, vp, vp, vp]
        L.vslam_cull_map_points_host.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, f64, vp, vp]
        L.vslam_frontend_run.argtypes = [vp, vp, i32, i32, vp, vp, f64, vp, vp, vp, vp, vp, vp]
        L.vslam_frontend_run_host.argtypes = [vp, vp, i32, f64, vp, vp, vp, vp, vp]
        L.vslam_orb_set_profiling.argtypes = [vp, i32]
        L.vslam_orb_read_profile.argtypes = [vp, vp, i32]
        L.vslam_launch_count.restype = C.c_ulonglong
        for name, args in _OPTIONAL.items():
            if hasattr(L, name):
                getattr(L, name).argtypes = args(vp, i32, f64)
        _LIB = L
    return _LIB


_OPTIONAL = {
    "vslam_orb_compute_host": lambda vp, i32, f64: [vp, vp, vp, vp, vp, vp, i32, vp, vp, vp],
    "vslam_bgr2gray_host": lambda vp, i32, f64: [vp, i32, i32, vp],
    "vslam_gftt_grid_host": lambda vp, i32, f64: [vp, i32, i32, i32, i32, i32, f64, f64, vp, vp, vp],
    "vslam_score_fundamental_host": lambda vp, i32, f64: [vp, vp, vp, i32, f64, vp, vp],
    "vslam_score_essential_host": lambda vp, i32, f64: [vp, vp, vp, vp, i32, f64, vp, vp],
    "vslam_ransac_fundamental_host": lambda vp, i32, f64: [vp, vp, i32, f64, f64, i32, C.c_uint64, vp, vp, vp],
    "vslam_ransac_essential_host": lambda vp, i32, f64: [vp, vp, i32, vp, f64, f64, i32, C.c_uint64, vp, vp, vp],
    "vslam_recover_pose_host": lambda vp, i32, f64: [vp, vp, vp, i32, vp, vp, vp, vp, vp],
    "vslam_triangulate_host": lambda vp, i32, f64: [vp, vp, vp, vp, i32, vp],
    "vslam_orb_debug_run_levels_host": lambda vp, i32, f64: [vp, vp, i32],
}


def check(rc):
    if rc != 0:
        raise VslamError("libvslam_b200 error %d: %s" % (rc, lib().vslam_last_error().decode()))


def ptr(a):
    """Host pointer of a C-contiguous numpy array (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def require_gpu():
    if lib().vslam_device_count() < 1:
        raise VslamError("no CUDA device visible: the B200 front end has no CPU fallback")


class OrbHandle:
    """Owns a vslam_orb workspace planned for one frame size."""

    def __init__(self, width, height, n_features=2000, scale_factor=1.2, n_levels=8, fast_threshold=7,
                 max_batch=1, kp_capacity=0, distribution="cv2", ini_threshold=20):
        """distribution: "cv2" = the reference's cv2.ORB retain-best by Harris response (default);
        "orbslam2" = per-cell ini/min FAST thresholds + quadtree (include/vslam_b200.h, VSLAM_DIST_ORBSLAM2)."""
        require_gpu()
        if distribution not in ("cv2", "orbslam2"):
            raise ValueError("unknown distribution %r" % (distribution,))
        self.params = OrbParams(int(width), int(height), int(n_features), int(n_levels), int(fast_threshold),
                                float(np.float32(scale_factor)), int(max_batch), int(kp_capacity))
        self._h = C.c_void_p()
        check(lib().vslam_orb_create(C.byref(self.params), C.byref(self._h)))
        if distribution == "orbslam2":
            check(lib().vslam_orb_set_distribution(self._h, 1, int(ini_threshold)))
        self.kp_capacity = lib().vslam_orb_kp_capacity(self._h)
        self.width, self.height, self.max_batch, self.n_levels = int(width), int(height), int(max_batch), int(n_levels)

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h:
            lib().vslam_orb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def level_info(self, level):
        w, h, p, q = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        s = C.c_float()
        check(lib().vslam_orb_level_info(self._h, level, C.byref(w), C.byref(h), C.byref(p), C.byref(s), C.byref(q)))
        return dict(w=w.value, h=h.value, pitch=p.value, scale=s.value, quota=q.value)

    def detect_and_compute_host(self, frames):
        """frames uint8 [B,H,W] -> (kp structured [B,cap], desc uint8 [B,cap,32], count int32 [B])."""
        frames = np.ascontiguousarray(frames, np.uint8)
        if frames.ndim == 2:
            frames = frames[None]
        b = frames.shape[0]
        assert frames.shape[1:] == (self.height, self.width), (frames.shape, self.height, self.width)
        kp = np.zeros((b, self.kp_capacity), KP_DTYPE)
        desc = np.zeros((b, self.kp_capacity, 32), np.uint8)
        cnt = np.zeros(b, np.int32)
        check(lib().vslam_orb_detect_and_compute_host(self._h, ptr(frames), b, ptr(kp), ptr(desc), ptr(cnt)))
        return kp, desc, cnt

    def detect_and_compute_host_bgr(self, frames_bgr):
        """frames uint8 [B,H,W,3] (BGR) -> like detect_and_compute_host; the gray conversion runs on the device."""
        frames_bgr = np.ascontiguousarray(frames_bgr, np.uint8)
        if frames_bgr.ndim == 3:
            frames_bgr = frames_bgr[None]
        b = frames_bgr.shape[0]
        assert frames_bgr.shape[1:] == (self.height, self.width, 3), frames_bgr.shape
        kp = np.zeros((b, self.kp_capacity), KP_DTYPE)
        desc = np.zeros((b, self.kp_capacity, 32), np.uint8)
        cnt = np.zeros(b, np.int32)
        check(lib().vslam_orb_detect_and_compute_host_bgr(self._h, ptr(frames_bgr), b, ptr(kp), ptr(desc), ptr(cnt)))
        return kp, desc, cnt

    def frontend_host(self, frames, ratio=0.75):
        frames = np.ascontiguousarray(frames, np.uint8)
        b = frames.shape[0]
        kp = np.zeros((b, self.kp_capacity), KP_DTYPE)
        desc = np.zeros((b, self.kp_capacity, 32), np.uint8)
        cnt = np.zeros(b, np.int32)
        match = np.zeros((max(b - 1, 1), self.kp_capacity, 3), np.int32)
        mcnt = np.zeros(max(b - 1, 1), np.int32)
        check(lib().vslam_frontend_run_host(self._h, ptr(frames), b, float(ratio), ptr(kp), ptr(desc), ptr(cnt),
                                            ptr(match), ptr(mcnt)))
        return kp, desc, cnt, match[:b - 1], mcnt[:b - 1]

    def debug_level(self, frame, level, blurred=False):
        info = self.level_info(level)
        out = np.zeros((info["h"], info["w"]), np.uint8)
        check(lib().vslam_orb_debug_level(self._h, frame, level, int(blurred), ptr(out)))
        return out

    def debug_candidates(self, frame, level):
        cap = 1 << 20
        x = np.zeros(cap, np.int32)
        y = np.zeros(cap, np.int32)
        s = np.zeros(cap, np.int32)
        n = C.c_int()
        check(lib().vslam_orb_debug_candidates(self._h, frame, level, cap, ptr(x), ptr(y), ptr(s), C.byref(n)))
        n = min(n.value, cap)
        return x[:n].copy(), y[:n].copy(), s[:n].copy()

    def debug_run_levels_host(self, frames):
        frames = np.ascontiguousarray(frames, np.uint8)
        if frames.ndim == 2:
            frames = frames[None]
        check(lib().vslam_orb_debug_run_levels_host(self._h, ptr(frames), frames.shape[0]))


def _desc32(d):
    """uint8 [n, 32] descriptors (an empty set may have any trailing shape); anything else is an error, not a reshape."""
    d = np.ascontiguousarray(d, np.uint8)
    if d.size == 0:
        return d.reshape(0, 32)
    if d.ndim != 2 or d.shape[1] != 32:
        raise ValueError("descriptors must be uint8 [n, 32] (256-bit ORB), got shape %r" % (d.shape,))
    return d


def bf_knn2_host(q, t):
    """uint8 [nq,32], [nt,32] -> idx int32 [nq,2], dist int32 [nq,2] (-1 = missing)."""
    require_gpu()
    q, t = _desc32(q), _desc32(t)
    idx = np.full((len(q), 2), -1, np.int32)
    dist = np.full((len(q), 2), -1, np.int32)
    check(lib().vslam_bf_knn2_hamming256_host(ptr(q), len(q), ptr(t), len(t), ptr(idx), ptr(dist)))
    return idx, dist


def bf_match_masked_host(q, t, mask=None):
    require_gpu()
    q, t = _desc32(q), _desc32(t)
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    if m is not None and m.shape != (len(q), len(t)):
        raise ValueError("mask shape %r does not match (%d, %d)" % (m.shape, len(q), len(t)))
    idx = np.full(len(q), -1, np.int32)
    dist = np.full(len(q), -1, np.int32)
    check(lib().vslam_bf_match_masked_host(ptr(q), len(q), ptr(t), len(t), ptr(m), ptr(idx), ptr(dist)))
    return idx, dist
