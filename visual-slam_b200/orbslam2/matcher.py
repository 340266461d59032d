"""Drop-in DescriptorMatcher (reference: src/orbslam2/matcher.py) on the B200 kernels.

Brute-force Hamming kNN-2, the masked k=1 search and the fundamental-matrix RANSAC run in
libvslam_b200.so; the Python-level filters keep the reference's exact arithmetic (Python
doubles on float32 values).  cv2 is only the container of `cv2.DMatch` and the drawer.
"""
import math

import numpy as np

from . import _native


def _cv2():
    import cv2
    return cv2


class DescriptorMatcher:
    """matcher.py:12-42."""

    def __init__(self, matcher_type='bruteforce-hamming', ratio_threshold=0.75):
        self.ratio_threshold = ratio_threshold
        if matcher_type == 'bruteforce-hamming':
            self.matcher_type = matcher_type
        elif matcher_type == 'flann':
            # matcher.py:30-40 builds an approximate LSH index; substituting brute force would
            # change results, so the drop-in refuses instead.
            raise NotImplementedError("matcher_type='flann' (approximate LSH) is outside the B200 hot path")
        else:
            raise ValueError(f"Unknown matcher type: {matcher_type}")

    # ------------------------------------------------------------------ matching
    def knn2(self, descriptors1, descriptors2):
        """Tensor-level kNN-2: (idx int32 [n, 2], dist int32 [n, 2]); -1 = no neighbour."""
        return _native.bf_knn2_host(descriptors1, descriptors2)

    def match(self, descriptors1, descriptors2, ratio_test=True):
        """matcher.py:44-83 -> list of cv2.DMatch in ascending queryIdx."""
        if descriptors1 is None or descriptors2 is None:
            return []
        if descriptors1.shape[0] == 0 or descriptors2.shape[0] == 0:
            return []
        if descriptors1.dtype != np.uint8:
            descriptors1 = np.uint8(descriptors1)
        if descriptors2.dtype != np.uint8:
            descriptors2 = np.uint8(descriptors2)
        idx, dist = _native.bf_knn2_host(descriptors1, descriptors2)
        d1 = dist[:, 0].astype(np.float32).astype(np.float64)
        d2 = dist[:, 1].astype(np.float32).astype(np.float64)
        two = idx[:, 1] >= 0
        one = (idx[:, 0] >= 0) & ~two
        if ratio_test:
            keep = (two & (d1 < self.ratio_threshold * d2)) | one      # m.distance < ratio * n.distance
        else:
            keep = two | one
        cv2 = _cv2()
        q = np.nonzero(keep)[0]
        # column lists + map(): the cv2.DMatch constructor is the floor; numpy scalar indexing per row doubled it
        return list(map(cv2.DMatch, q.tolist(), idx[q, 0].tolist(), [0] * len(q), dist[q, 0].astype(np.float32).tolist()))

    def match_with_mask(self, descriptors1, descriptors2, mask):
        """matcher.py:85-107: k = 1 restricted by mask, sorted by distance (stable)."""
        if descriptors1 is None or descriptors2 is None:
            return []
        d1 = np.ascontiguousarray(descriptors1, np.uint8)
        d2 = np.ascontiguousarray(descriptors2, np.uint8)
        if d1.shape[0] == 0 or d2.shape[0] == 0:
            return []
        if mask is not None:
            mask = np.asarray(mask)
            if mask.shape != (d1.shape[0], d2.shape[0]):       # cv2 asserts on this; the C side would read out of bounds
                raise ValueError("mask shape %r does not match (%d, %d)" % (mask.shape, d1.shape[0], d2.shape[0]))
        idx, dist = _native.bf_match_masked_host(d1, d2, mask)
        cv2 = _cv2()
        matches = [cv2.DMatch(int(q), int(idx[q]), 0, float(dist[q])) for q in range(len(idx)) if idx[q] >= 0]
        return sorted(matches, key=lambda x: x.distance)

    # ------------------------------------------------------------------- filters
    def filter_matches_by_geometric_distance(self, keypoints1, keypoints2, matches, threshold_percent, image_shape):
        """matcher.py:109-142 (Python doubles on float32 coordinates, as the reference)."""
        height, width = image_shape[:2]
        max_dist = ((width + height) / 2.0) * threshold_percent
        filtered = []
        for m in matches:
            pt1 = keypoints1[m.queryIdx].pt
            pt2 = keypoints2[m.trainIdx].pt
            if math.hypot(pt2[0] - pt1[0], pt2[1] - pt1[1]) <= max_dist:
                filtered.append(m)
        return filtered

    def filter_matches_by_distance(self, matches, distance_threshold=None):
        """matcher.py:144-169: stable sort by distance, keep < 2 x median."""
        if not matches:
            return []
        matches = sorted(matches, key=lambda x: x.distance)
        if distance_threshold is None:
            distance_threshold = np.median([m.distance for m in matches]) * 2.0
        return [m for m in matches if m.distance < distance_threshold]

    def filter_matches_by_fundamental(self, keypoints1, keypoints2, matches, threshold=3.0):
        """matcher.py:171-200: RANSAC fundamental matrix (conf 0.99) -> (matches, bool mask)."""
        if len(matches) < 8:
            return matches, np.ones(len(matches), dtype=bool)
        points1 = np.float32([keypoints1[m.queryIdx].pt for m in matches])
        points2 = np.float32([keypoints2[m.trainIdx].pt for m in matches])
        from .utils import find_fundamental_ransac
        F, mask = find_fundamental_ransac(points1, points2, threshold, 0.99)
        mask = mask.ravel().astype(bool)
        return [m for m, inl in zip(matches, mask) if inl], mask

    def draw_matches(self, img1, keypoints1, img2, keypoints2, matches, flags=0):
        """matcher.py:202-217 (CPU visualisation, delegated to cv2 like the reference)."""
        return _cv2().drawMatches(img1, keypoints1, img2, keypoints2, matches, None, flags=flags)
