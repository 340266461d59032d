"""CPU oracle of LocalMapper._cull_map_points' per-point test.

TEST INFRASTRUCTURE ONLY (imported by tests/ and nothing else).

Restates /root/reference/src/orbslam2/local_mapper.py:205-237 on flat arrays: a map point survives when it has
at least `min_observations` observations and, walking them in stored order, none has a reprojection error
above 5 px.  P = K [R | t] per keyframe (utils.py:137-159), proj = P [X 1], error = |proj_xy / proj_z - kp.pt|
in float64; the reference writes `err > 5.0`, so a NaN error (point on the principal plane) does not reject.
Pinned against tests/golden/cull.npz (made by the reference's own LocalMapper, make_cull_golden.py).
"""
import numpy as np


def projection_matrices(K, poses):
    """[n_kf, 3, 4] float64, camera_matrix @ [R | t] like utils.compute_projection_matrix."""
    poses = np.asarray(poses, np.float64).reshape(-1, 4, 4)
    return np.array([np.asarray(K, np.float64) @ np.hstack((T[:3, :3], T[:3, 3].reshape(3, 1))) for T in poses]).reshape(-1, 3, 4)


def cull(points, offs, obs_kf, obs_xy, P, min_observations=2, max_error=5.0):
    """-> (valid bool[n], err float64[n]): err = the largest error among the observations the reference
    looks at (it stops at the first one above the gate); 0 for a point it rejects on the count alone."""
    points = np.asarray(points, np.float64).reshape(-1, 3)
    n = len(points)
    valid, err = np.zeros(n, bool), np.zeros(n, np.float64)
    for i in range(n):
        lo, hi = int(offs[i]), int(offs[i + 1])
        if hi - lo < min_observations:
            continue
        ok, worst = True, 0.0
        with np.errstate(all="ignore"):
            for j in range(lo, hi):
                proj = P[int(obs_kf[j])] @ np.append(points[i], 1.0)
                proj = proj[:2] / proj[2]
                e = np.linalg.norm(proj - np.asarray(obs_xy[j], np.float64))
                if e > worst:                   # a NaN error neither rejects nor is reported
                    worst = e
                if e > max_error:
                    ok = False
                    break
        valid[i], err[i] = ok, worst
    return valid, err
