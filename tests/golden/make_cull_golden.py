"""Generates tests/golden/cull.npz: which map points the reference's own LocalMapper._cull_map_points
(/root/reference/src/orbslam2/local_mapper.py:205-245) keeps, on seeded synthetic maps.  Run in the build container:
    python tests/golden/make_cull_golden.py
Each case stores the map as flat arrays (points, CSR observation lists, keypoint coordinates, keyframe poses)
plus the ids that survived, so the tests need neither the reference nor cv2.KeyPoint objects.
"""
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, "/root/reference/src")
from orbslam2.local_mapper import LocalMapper   # the reference's own class


def rodrigues(v):
    return cv2.Rodrigues(np.asarray(v, np.float64))[0]


def make_case(rng, n_kf, n_pts, noise, min_obs, drop_single):
    K = np.array([[520.0, 0, 320.5], [0, 515.0, 240.25], [0, 0, 1]])
    poses = []
    for k in range(n_kf):
        T = np.eye(4)
        T[:3, :3] = rodrigues(rng.normal(0, 0.05, 3))
        T[:3, 3] = rng.normal(0, 0.3, 3) + [0.2 * k, 0, 0]
        poses.append(T)
    pts = np.c_[rng.uniform(-4, 4, n_pts), rng.uniform(-3, 3, n_pts), rng.uniform(2, 12, n_pts)]
    kp_xy = [[] for _ in range(n_kf)]
    offs, obs_kf, obs_kp = [0], [], []
    for i in range(n_pts):
        if drop_single and i % 7 == 0:
            seen = rng.choice(n_kf, size=1, replace=False)          # fewer than min_observations
        elif i % 11 == 0:
            seen = np.array([], int)                                  # no observation at all
        else:
            seen = np.sort(rng.choice(n_kf, size=int(rng.integers(2, n_kf + 1)), replace=False))
            if i % 3 == 0:
                seen = seen[::-1]                                     # dict order is insertion order, not sorted
        for k in seen:
            T = poses[k]
            c = T[:3, :3] @ pts[i] + T[:3, 3]
            uv = (K @ c)[:2] / c[2]
            # most observations close to the projection, some near and beyond the 5 px gate
            r = rng.choice([0.3, 3.0, 4.9, 5.1, 9.0], p=[0.6, 0.15, 0.1, 0.1, 0.05]) if noise else 0.0
            a = rng.uniform(0, 2 * np.pi)
            uv = (uv + r * np.array([np.cos(a), np.sin(a)])).astype(np.float32)    # cv2.KeyPoint.pt is float32
            obs_kf.append(int(k)); obs_kp.append(len(kp_xy[k])); kp_xy[k].append(uv)
        offs.append(len(obs_kf))
    if n_pts > 5:                                                     # a point on a camera's principal plane (w = 0)
        T = poses[0]
        pts[5] = np.linalg.inv(T)[:3, :3] @ np.array([0.5, 0.2, 0.0]) + np.linalg.inv(T)[:3, 3]
    return K, poses, pts, kp_xy, np.array(offs, np.int32), np.array(obs_kf, np.int32), np.array(obs_kp, np.int32)


def run_reference(K, poses, pts, kp_xy, offs, obs_kf, obs_kp, min_obs):
    lm = LocalMapper(K)
    lm.min_observations = min_obs
    for k, T in enumerate(poses):
        kps = [cv2.KeyPoint(float(x), float(y), 31.0) for x, y in kp_xy[k]]
        lm.keyframes.append(dict(id=k, image=None, keypoints=kps, descriptors=None, pose=T, map_points=[]))
    for i in range(len(pts)):
        obs = {int(obs_kf[j]): int(obs_kp[j]) for j in range(offs[i], offs[i + 1])}
        lm.map_points.append(dict(id=i, position=pts[i].copy(), color=None, observed_keyframes=obs, descriptor=None))
    with np.errstate(all="ignore"):
        lm._cull_map_points()
    kept = np.array([mp["id"] for mp in lm.map_points], np.int32)
    per_kf = [np.array(kf["map_points"], np.int32) for kf in lm.keyframes]
    return kept, per_kf


rng = np.random.default_rng(77)
out = {}
cases = [(3, 40, True, 2, True), (6, 600, True, 2, True), (5, 300, True, 3, False), (4, 200, False, 2, False), (2, 0, True, 2, False)]
for n, (n_kf, n_pts, noise, min_obs, drop) in enumerate(cases):
    K, poses, pts, kp_xy, offs, obs_kf, obs_kp = make_case(rng, n_kf, n_pts, noise, min_obs, drop)
    kept, per_kf = run_reference(K, poses, pts, kp_xy, offs, obs_kf, obs_kp, min_obs)
    p = "c%d_" % n
    out[p + "K"], out[p + "poses"], out[p + "points"] = K, np.array(poses), pts.reshape(-1, 3)
    out[p + "offs"], out[p + "obs_kf"], out[p + "obs_kp"] = offs, obs_kf, obs_kp
    out[p + "kp_off"] = np.cumsum([0] + [len(k) for k in kp_xy]).astype(np.int32)
    out[p + "kp_xy"] = np.array([xy for k in kp_xy for xy in k], np.float32).reshape(-1, 2)
    out[p + "min_obs"], out[p + "kept"] = np.array(min_obs), kept
    out[p + "kf_lists_off"] = np.cumsum([0] + [len(a) for a in per_kf]).astype(np.int32)
    out[p + "kf_lists"] = np.concatenate(per_kf) if per_kf else np.zeros(0, np.int32)
    print("case", n, "points", n_pts, "kept", len(kept))
out["n"] = np.array(len(cases))
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cull.npz"), **out)
print("wrote cull.npz")
