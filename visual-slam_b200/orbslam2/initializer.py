"""Drop-in MapInitializer (reference: src/orbslam2/initializer.py) on the B200 kernels.

Same constructor, public attributes, methods and return tuples as the reference class; the
essential-matrix RANSAC, pose recovery and triangulation run in libvslam_b200.so.  The
reference's debug prints (initializer.py:71,98-102,124) are not part of the contract and are
omitted; pass verbose=True to get them back.
"""
import numpy as np

from .utils import (calculate_essential_matrix, compute_projection_matrix, convert_to_3d_points, recover_pose,
                    triangulate_points)


class MapInitializer:
    """initializer.py:12-33."""

    def __init__(self, camera_matrix, min_matches=10, min_inliers_ratio=0.9, verbose=False):
        self.camera_matrix = camera_matrix
        self.min_matches = min_matches
        self.min_inliers_ratio = min_inliers_ratio     # stored and unused, as in the reference
        self.initialization_done = False
        self.first_frame_keypoints = None
        self.first_frame_descriptors = None
        self.first_frame_image = None
        self.current_frame_keypoints = None
        self.verbose = verbose

    def set_first_frame(self, keypoints, descriptors, image):
        """initializer.py:35-47."""
        self.first_frame_keypoints = keypoints
        self.first_frame_descriptors = descriptors
        self.first_frame_image = image.copy()
        self.initialization_done = False

    def initialize(self, current_keypoints, current_descriptors, matcher, current_image):
        """initializer.py:49-152 -> (success, R, t, initial_map_points, matches)."""
        if self.first_frame_keypoints is None or self.first_frame_descriptors is None:
            return False, None, None, None, None
        matches = matcher.match(self.first_frame_descriptors, current_descriptors)
        if len(matches) < self.min_matches:
            if self.verbose:
                print(f"Not enough matches for initialization: {len(matches)} < {self.min_matches}")
            return False, None, None, None, matches
        points1 = np.float32([self.first_frame_keypoints[m.queryIdx].pt for m in matches])
        points2 = np.float32([current_keypoints[m.trainIdx].pt for m in matches])
        E, mask = calculate_essential_matrix(points1, points2, self.camera_matrix, threshold=3.0)
        if E is None:
            return False, None, None, None, matches
        _, R, t, mask_pose, *_ = recover_pose(E, points1, points2, self.camera_matrix, mask)
        t = t.reshape(3, 1) if t.ndim == 1 else t
        P1 = compute_projection_matrix(np.eye(3), np.zeros((3, 1)), self.camera_matrix)
        P2 = compute_projection_matrix(R, t, self.camera_matrix)
        valid_matches = [m for m, valid in zip(matches, mask_pose.ravel().astype(bool)) if valid]
        if not valid_matches:
            return False, None, None, None, valid_matches
        points1 = np.float32([self.first_frame_keypoints[m.queryIdx].pt for m in valid_matches])
        points2 = np.float32([current_keypoints[m.trainIdx].pt for m in valid_matches])
        points_4d = triangulate_points(points1, points2, P1, P2)
        points_3d = convert_to_3d_points(points_4d)
        if self.verbose:
            print("points_4d", points_4d.shape, "points_3d", points_3d.shape)
        # points behind either camera are dropped (initializer.py:105-120)
        depth1 = points_3d[:, 2]
        depth2 = (points_3d.astype(np.float64) @ R.T + t.ravel())[:, 2]
        valid_indices = np.nonzero((depth1 > 0) & (depth2 > 0))[0].tolist()
        points_3d = points_3d[valid_indices]
        valid_matches = [valid_matches[i] for i in valid_indices]
        if len(points_3d) < self.min_matches // 2:
            if self.verbose:
                print(f"Insufficient valid 3D points after filtering: {len(points_3d)}")
            return False, None, None, None, valid_matches
        colors = []
        img = self.first_frame_image
        for match in valid_matches:
            kp = self.first_frame_keypoints[match.queryIdx]
            x, y = int(kp.pt[0]), int(kp.pt[1])
            if 0 <= x < img.shape[1] and 0 <= y < img.shape[0]:
                color = img[y, x, :] if img.ndim == 3 else np.array([img[y, x]] * 3)
            else:
                color = np.array([0, 0, 255])
            colors.append(color)
        initial_map_points = [{
            'position': pt,
            'color': colors[i],
            'keypoint_references': {0: m.queryIdx, 1: m.trainIdx},
            'observed_frames': [0, 1]
        } for i, (pt, m) in enumerate(zip(points_3d, valid_matches))]
        self.initialization_done = True
        self.current_frame_keypoints = current_keypoints
        return True, R, t, initial_map_points, valid_matches

    def draw_initialization(self, first_image, current_image, matches):
        """initializer.py:154-181 (CPU visualisation, delegated to cv2 like the reference)."""
        import cv2
        first_img = cv2.cvtColor(first_image, cv2.COLOR_GRAY2BGR) if len(first_image.shape) == 2 else first_image.copy()
        current_img = cv2.cvtColor(current_image, cv2.COLOR_GRAY2BGR) if len(current_image.shape) == 2 else current_image.copy()
        return cv2.drawMatches(first_img, self.first_frame_keypoints, current_img, self.current_frame_keypoints, matches,
                               None, flags=cv2.DrawMatchesFlags_NOT_DRAW_SINGLE_POINTS, matchColor=(0, 255, 0),
                               singlePointColor=(255, 0, 0))
