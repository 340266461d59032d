"""Drop-in ORBExtractor (reference: src/orbslam2/extractor.py) on the B200 kernels.

Same constructor, attributes, method names, argument meaning and return types as the
reference class; every cv2 compute call of the reference is replaced by a call into
libvslam_b200.so (CUDA, sm_100a).  cv2 is used only as the container of the
`cv2.KeyPoint` return type and for drawing (extractor.py:162-174), never for compute.
There is no CPU fallback: without the CUDA library / a GPU the methods raise.
"""
import numpy as np

from . import _native


def _cv2():
    import cv2
    return cv2


class ORBExtractor:
    """ORB feature extractor (extractor.py:13-48)."""

    def __init__(self, n_features=2000, scale_factor=1.2, n_levels=8, ini_threshold=20, min_threshold=7,
                 distribution="cv2"):
        """The five reference parameters (extractor.py:19-35), plus one opt-in that the reference does not
        have: distribution="orbslam2" makes `detect_and_compute` spread keypoints the ORB-SLAM2 way (FAST at
        ini_threshold per 30-px cell with fallback to min_threshold, quadtree to the level quota, response =
        FAST score) instead of cv2.ORB's Harris retain-best.  The default reproduces the reference."""
        if distribution not in ("cv2", "orbslam2"):
            raise ValueError("distribution must be 'cv2' or 'orbslam2'")
        self.distribution = distribution
        self.n_features = n_features
        self.scale_factor = scale_factor
        self.n_levels = n_levels
        self.ini_threshold = ini_threshold      # stored and unused by the reference (extractor.py:34); used by "orbslam2"
        self.min_threshold = min_threshold      # reaches cv2.ORB_create as fastThreshold (extractor.py:47)
        self._handles = {}

    # ------------------------------------------------------------------ helpers
    def _handle(self, w, h, n_features=None):
        key = (w, h, n_features or self.n_features)
        hd = self._handles.get(key)
        if hd is None:
            if len(self._handles) >= 4:          # bound the workspace cache
                self._handles.pop(next(iter(self._handles))).close()
            hd = _native.OrbHandle(w, h, n_features=key[2], scale_factor=self.scale_factor,
                                   n_levels=self.n_levels, fast_threshold=self.min_threshold, max_batch=1,
                                   distribution=self.distribution, ini_threshold=self.ini_threshold)
            self._handles[key] = hd
        return hd

    @staticmethod
    def _gray(image):
        """extractor.py:61-62: cvtColor(BGR2GRAY) when the image has a channel axis."""
        image = np.asarray(image)
        if image.ndim > 2:
            if image.shape[2] == 1:
                return np.ascontiguousarray(image[:, :, 0], np.uint8)
            bgr = np.ascontiguousarray(image[:, :, :3], np.uint8)
            gray = np.empty(bgr.shape[:2], np.uint8)
            _native.require_gpu()
            _native.check(_native.lib().vslam_bgr2gray_host(_native.ptr(bgr), bgr.shape[0], bgr.shape[1],
                                                            _native.ptr(gray)))
            return gray
        return np.ascontiguousarray(image, np.uint8)

    @staticmethod
    def _keypoints(kp, n):
        """structured rows -> tuple of cv2.KeyPoint.  Column lists + map(): the constructor calls are the floor
        (~0.35 us each); per-row field access on the structured array cost five times that."""
        cv2 = _cv2()
        k = kp[:n]
        return tuple(map(cv2.KeyPoint, k["x"].tolist(), k["y"].tolist(), k["size"].tolist(), k["angle"].tolist(),
                         k["response"].tolist(), k["octave"].tolist()))

    # ---------------------------------------------------------------------- API
    def detect_and_compute(self, image):
        """extractor.py:50-67 -> (tuple of cv2.KeyPoint, uint8 [n, 32]) or ((), None)."""
        image = np.asarray(image)
        if image.ndim == 3 and image.shape[2] == 3 and image.dtype == np.uint8:
            # colour frame: one upload, cvtColor(BGR2GRAY) runs on the device in front of the pyramid
            h, w = image.shape[:2]
            hd = self._handle(w, h)
            kp, desc, cnt = hd.detect_and_compute_host_bgr(image)
            gray = None
        else:
            gray = self._gray(image)
            h, w = gray.shape
            hd = self._handle(w, h)
            kp, desc, cnt = hd.detect_and_compute_host(gray)
        n = int(cnt[0])
        if n > hd.kp_capacity:
            if gray is None:
                gray = self._gray(image)
            # retain-best kept more ties than the default slack: re-run with room for all of them
            big = _native.OrbHandle(w, h, n_features=self.n_features, scale_factor=self.scale_factor,
                                    n_levels=self.n_levels, fast_threshold=self.min_threshold, max_batch=1,
                                    kp_capacity=n, distribution=self.distribution,
                                    ini_threshold=self.ini_threshold)
            try:
                kp, desc, cnt = big.detect_and_compute_host(gray)
            finally:
                big.close()
            n = int(cnt[0])
        if n == 0:
            return (), None
        return self._keypoints(kp[0], n), np.ascontiguousarray(desc[0, :n])

    def detect_and_compute_arrays(self, image):
        """Tensor-level variant without the cv2.KeyPoint objects: (structured kp [n], desc [n, 32])."""
        gray = self._gray(image)
        h, w = gray.shape
        kp, desc, cnt = self._handle(w, h).detect_and_compute_host(gray)
        n = min(int(cnt[0]), kp.shape[1])
        return kp[0, :n].copy(), desc[0, :n].copy()

    def _compute_arrays(self, gray, keypoints):
        h, w = gray.shape
        n = len(keypoints)
        if n == 0:
            return np.zeros((0, 32), np.uint8), np.zeros(0, np.int32)
        x = np.array([k.pt[0] for k in keypoints], np.float32)
        y = np.array([k.pt[1] for k in keypoints], np.float32)
        a = np.array([k.angle for k in keypoints], np.float32)
        o = np.array([k.octave for k in keypoints], np.int32)
        desc = np.zeros((n, 32), np.uint8)
        keep = np.zeros(n, np.int32)
        m = _native.C.c_int()
        hd = self._handle(w, h)
        _native.check(_native.lib().vslam_orb_compute_host(hd.handle, _native.ptr(gray), _native.ptr(x), _native.ptr(y),
                                                           _native.ptr(a), _native.ptr(o), n, _native.ptr(desc),
                                                           _native.ptr(keep), _native.C.byref(m)))
        return desc[:m.value].copy(), keep[:m.value].copy()

    def compute(self, image, keypoints, aligned=False):
        """extractor.py:68-83.  Returns the CALLER'S keypoint list (unfiltered) with the
        descriptors of the in-border keypoints only, exactly like the reference (which
        discards the filtered list cv2 returns).  aligned=True is an opt-in deviation that
        returns the filtered keypoints instead."""
        gray = self._gray(image)
        desc, keep = self._compute_arrays(gray, keypoints)
        if len(desc) == 0:
            desc = None
        if aligned:
            return [keypoints[i] for i in keep], desc
        return keypoints, desc

    def distribute_keypoints(self, image, n_features=None, aligned=False):
        """extractor.py:85-144: 8x8 grid of goodFeaturesToTrack + unsteered BRIEF at level 0."""
        if n_features is None:
            n_features = self.n_features
        gray = self._gray(image)
        h, w = gray.shape
        rows = cols = 8
        per_cell = n_features // (rows * cols)
        cv2 = _cv2()
        all_keypoints = []
        if per_cell <= 0:
            # cv2.goodFeaturesToTrack(maxCorners=0) means "no limit" (extractor.py:123 with n_features < 64): every
            # corner above the quality level at >= 10 px from the accepted ones.  A cell cannot hold more than
            # (cw / 10 + 1) * (ch / 10 + 1) such corners; the kernel's per-cell bound is 1024.
            bound = (w // cols // 10 + 2) * (h // rows // 10 + 2)
            if bound > 1024:
                raise NotImplementedError("distribute_keypoints with n_features < 64 on cells larger than ~300x300 px")
            per_cell = bound
        if per_cell > 0:
            xy = np.zeros((rows * cols * per_cell, 2), np.float32)
            cc = np.zeros(rows * cols, np.int32)
            n = _native.C.c_int()
            _native.require_gpu()
            _native.check(_native.lib().vslam_gftt_grid_host(_native.ptr(gray), h, w, rows, cols, per_cell, 0.01, 10.0,
                                                             _native.ptr(xy), _native.ptr(cc), _native.C.byref(n)))
            all_keypoints = [cv2.KeyPoint(float(px), float(py), 31) for px, py in xy[:n.value]]
        if all_keypoints:
            desc, keep = self._compute_arrays(gray, all_keypoints)
            descriptors = desc if len(desc) else None
            if aligned:
                return [all_keypoints[i] for i in keep], descriptors
        else:
            descriptors = None
        return all_keypoints, descriptors

    def extract_features(self, image, distributed=True):
        """extractor.py:146-160."""
        if distributed:
            return self.distribute_keypoints(image)
        return self.detect_and_compute(image)

    def draw_keypoints(self, image, keypoints):
        """extractor.py:162-174 (CPU visualisation, delegated to cv2 like the reference)."""
        cv2 = _cv2()
        return cv2.drawKeypoints(image, keypoints, None, flags=cv2.DRAW_MATCHES_FLAGS_DRAW_RICH_KEYPOINTS)
