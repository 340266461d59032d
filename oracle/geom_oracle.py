"""CPU restatement (numpy) of the geometry stages of the hot path.

TEST INFRASTRUCTURE ONLY (same rules as oracle/oracle.py): imported by tests/ and by
__graft_entry__.smoke(), never by the product package.

The reference reaches this arithmetic through OpenCV calib3d (third-party, not vendored;
opencv-python-headless 4.13.0.92 is the de-facto pin):
    matcher.py:192 / local_mapper.py:136   cv2.findFundamentalMat(p1, p2, FM_RANSAC, thr, 0.99)
    utils.py:124-125 / tracker.py:242      cv2.findEssentialMat(p1, p2, K, RANSAC, prob, thr)
    utils.py:133 / tracker.py:249          cv2.recoverPose(E, p1, p2, K, mask)
    utils.py:59                            cv2.triangulatePoints(P1, P2, p1.T, p2.T)
Restated here following OpenCV's published algorithm:
  * the two error functions that define the inlier masks (exact: pinned by
    tests/test_oracle_golden.py against the masks cv2 returned for its own F / E),
  * decomposeEssentialMat + the four-candidate cheirality test of recoverPose,
  * the per-point DLT of triangulatePoints.
The RANSAC loops themselves are NOT restated: OpenCV's hypothesis stream depends on its RNG
and SVD internals, so the estimators are pinned by the golden F / E / masks generated from
the reference (tests/golden/make_golden.py) and compared by inlier-set agreement.
"""
import numpy as np


def fundamental_error(F, p1, p2):
    """FMEstimatorCallback::computeError: max(d1^2, d2^2) in double, returned as float32."""
    F = np.asarray(F, np.float64).reshape(9)
    x1, y1 = p1[:, 0].astype(np.float64), p1[:, 1].astype(np.float64)
    x2, y2 = p2[:, 0].astype(np.float64), p2[:, 1].astype(np.float64)
    a = F[0] * x1 + F[1] * y1 + F[2]
    b = F[3] * x1 + F[4] * y1 + F[5]
    c = F[6] * x1 + F[7] * y1 + F[8]
    s2 = 1.0 / (a * a + b * b)
    d2 = x2 * a + y2 * b + c
    a = F[0] * x2 + F[3] * y2 + F[6]
    b = F[1] * x2 + F[4] * y2 + F[7]
    c = F[2] * x2 + F[5] * y2 + F[8]
    s1 = 1.0 / (a * a + b * b)
    d1 = x1 * a + y1 * b + c
    return np.maximum(d1 * d1 * s1, d2 * d2 * s2).astype(np.float32)


def score_fundamental(F, p1, p2, thr):
    """Inlier mask cv2.findFundamentalMat(FM_RANSAC) reports for its returned F."""
    return (fundamental_error(F, p1, p2) <= np.float32(thr * thr)).astype(np.uint8)


def normalise(p, K):
    K = np.asarray(K, np.float64)
    q = np.asarray(p, np.float64).copy()
    q[:, 0] = (q[:, 0] - K[0, 2]) / K[0, 0]
    q[:, 1] = (q[:, 1] - K[1, 2]) / K[1, 1]
    return q


def essential_error(E, q1, q2):
    """EMEstimatorCallback::computeError (Sampson) on normalised double points -> float32."""
    E = np.asarray(E, np.float64).reshape(9)
    x1, y1, x2, y2 = q1[:, 0], q1[:, 1], q2[:, 0], q2[:, 1]
    Ex0 = E[0] * x1 + E[1] * y1 + E[2]
    Ex1 = E[3] * x1 + E[4] * y1 + E[5]
    Ex2 = E[6] * x1 + E[7] * y1 + E[8]
    Et0 = E[0] * x2 + E[3] * y2 + E[6]
    Et1 = E[1] * x2 + E[4] * y2 + E[7]
    x2tEx1 = x2 * Ex0 + y2 * Ex1 + Ex2
    return (x2tEx1 * x2tEx1 / (Ex0 * Ex0 + Ex1 * Ex1 + Et0 * Et0 + Et1 * Et1)).astype(np.float32)


def score_essential(E, p1, p2, K, thr):
    K = np.asarray(K, np.float64)
    t = thr / ((K[0, 0] + K[1, 1]) / 2)
    return (essential_error(E, normalise(p1, K), normalise(p2, K)) <= np.float32(t * t)).astype(np.uint8)


def triangulate(P1, P2, p1, p2):
    """cv2.triangulatePoints: per-point 4x4 DLT, null vector by SVD -> float32 [4, N]."""
    return triangulate64(P1, P2, p1, p2).astype(np.float32)


def triangulate64(P1, P2, p1, p2):
    P1, P2 = np.asarray(P1, np.float64), np.asarray(P2, np.float64)
    p1, p2 = np.asarray(p1, np.float64), np.asarray(p2, np.float64)
    A = np.stack([p1[:, 0, None] * P1[2] - P1[0], p1[:, 1, None] * P1[2] - P1[1],
                  p2[:, 0, None] * P2[2] - P2[0], p2[:, 1, None] * P2[2] - P2[1]], axis=1)
    _, _, Vt = np.linalg.svd(A)
    return Vt[:, 3, :].T


def decompose_essential(E):
    U, _, Vt = np.linalg.svd(np.asarray(E, np.float64))
    if np.linalg.det(U) < 0:
        U = -U
    if np.linalg.det(Vt) < 0:
        Vt = -Vt
    W = np.array([[0.0, 1, 0], [-1, 0, 0], [0, 0, 1]])
    return U @ W @ Vt, U @ W.T @ Vt, U[:, 2].copy()


def recover_pose(E, p1, p2, K, mask=None, dist=50.0):
    """cv2.recoverPose -> (n_good, R, t [3,1], mask uint8 [n])."""
    R1, R2, t = decompose_essential(E)
    q1, q2 = normalise(p1, K), normalise(p2, K)
    P0 = np.hstack([np.eye(3), np.zeros((3, 1))])
    cands = [(R1, t), (R2, t), (R1, -t), (R2, -t)]
    masks = []
    for R, tt in cands:
        P = np.hstack([R, tt.reshape(3, 1)])
        Q = triangulate64(P0, P, q1, q2)
        with np.errstate(divide="ignore", invalid="ignore"):
            m = Q[2] * Q[3] > 0
            Qn = Q / Q[3]
            m &= Qn[2] < dist
            z2 = (P @ Qn)[2]
            m &= (z2 > 0) & (z2 < dist)
        if mask is not None:
            m &= np.asarray(mask).reshape(-1) != 0
        masks.append(m)
    good = [int(m.sum()) for m in masks]
    if good[0] >= good[1] and good[0] >= good[2] and good[0] >= good[3]:
        w = 0
    elif good[1] >= good[0] and good[1] >= good[2] and good[1] >= good[3]:
        w = 1
    elif good[2] >= good[0] and good[2] >= good[1] and good[2] >= good[3]:
        w = 2
    else:
        w = 3
    return good[w], cands[w][0], cands[w][1].reshape(3, 1), masks[w].astype(np.uint8)
