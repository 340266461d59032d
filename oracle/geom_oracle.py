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
  * the RANSAC / LMedS loops themselves (RANSACPointSetRegistrator::run): OpenCV's RNG stream (seeded -1 on
    every call), subset drawing with duplicate and collinearity rejection, every model of a subset scored,
    RANSACUpdateNumIters.  Pinned: on the golden pairs the replay returns cv2's own masks, and the 7 points
    cv2's returned F interpolates are a row of the generated subset table
    (oracle/tools/probe_cv2_ransac_stream.py, tests/test_oracle_golden.py).  What is not reproducible is
    OpenCV's rounding inside its minimal solvers: the solvers here return every real solution polished to
    machine precision, and cv2 agrees with that wherever its own roots are accurate (all but one golden case:
    pair12 at threshold 1, where cv2's winning root violates the cubic constraints by 5e-5).
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


# ---------------------------------------------------------------------------------------------
# OpenCV's RANSAC loop (RANSACPointSetRegistrator::run), restated from its published algorithm so
# that the estimators can be compared with the reference's outputs sample for sample.
# cv2.findFundamentalMat / findEssentialMat are deterministic: every call seeds RNG(-1).

class CvRNG:
    """cv::RNG: multiply-with-carry, state = (uint32)state * 4164903690 + (state >> 32)."""
    COEFF = 4164903690

    def __init__(self, state=0xFFFFFFFFFFFFFFFF):
        self.state = state if state else 0xFFFFFFFF

    def next(self):
        self.state = ((self.state & 0xFFFFFFFF) * self.COEFF + (self.state >> 32)) & 0xFFFFFFFFFFFFFFFF
        return self.state & 0xFFFFFFFF

    def uniform(self, a, b):
        return a if a == b else int(self.next() % (b - a) + a)


def _last_point_collinear(pts):
    """haveCollinearPoints(m, count): only the LAST point is tested against the lines through pairs of
    earlier ones (float32 differences widened to double, as `double dx1 = ptr[j].x - ptr[i].x`)."""
    pts = np.asarray(pts, np.float32)
    i = len(pts) - 1
    eps = float(np.finfo(np.float32).eps)
    for j in range(i):
        dx1 = float(pts[j][0] - pts[i][0])
        dy1 = float(pts[j][1] - pts[i][1])
        for k in range(j):
            dx2 = float(pts[k][0] - pts[i][0])
            dy2 = float(pts[k][1] - pts[i][1])
            if abs(dx2 * dy1 - dy2 * dx1) <= eps * (abs(dx1) + abs(dy1) + abs(dx2) + abs(dy2)):
                return True
    return False


def cv2_sample_table(p1, p2, model_points, n_rows, check_collinear, max_attempts=10000):
    """The first n_rows subsets RANSACPointSetRegistrator::getSubset draws for `count` points: per
    subset, model_points draws of rng.uniform(0, count) with duplicate rejection; when check_collinear
    (FMEstimatorCallback::checkSubset) a subset whose last point is collinear with two earlier ones in
    either image is redrawn (up to 10000 attempts)."""
    rng = CvRNG()
    count = len(p1)
    rows = []
    for _ in range(n_rows):
        for _attempt in range(max_attempts):
            idx = []
            for i in range(model_points):
                v = rng.uniform(0, count)
                while v in idx:
                    v = rng.uniform(0, count)
                idx.append(v)
            if not check_collinear or not (_last_point_collinear(p1[idx]) or _last_point_collinear(p2[idx])):
                break
        else:
            return rows
        rows.append(idx)
    return rows


def ransac_update_num_iters(p, ep, model_points, max_iters):
    """cv::RANSACUpdateNumIters."""
    p = min(max(p, 0.0), 1.0)
    ep = min(max(ep, 0.0), 1.0)
    num = max(1.0 - p, np.finfo(np.float64).tiny)
    denom = 1.0 - (1.0 - ep) ** model_points
    if denom < np.finfo(np.float64).tiny:
        return 0
    num, denom = np.log(num), np.log(denom)
    if denom >= 0 or -num >= max_iters * (-denom):
        return max_iters
    return int(np.rint(num / denom))


def solve_cubic(c):
    """cv::solveCubic for c[0] x^3 + c[1] x^2 + c[2] x + c[3] = 0 -> real roots in OpenCV's order."""
    a0, a1, a2, a3 = (float(v) for v in c)
    if a0 == 0:
        if a1 == 0:
            return [] if a2 == 0 else [-a3 / a2]
        d = a2 * a2 - 4 * a1 * a3
        if d < 0:
            return []
        d = np.sqrt(d)
        q1, q2 = (-a2 + d) * 0.5, (a2 + d) * -0.5
        r = [q1 / a1, a3 / q1] if abs(q1) > abs(q2) else [q2 / a1, a3 / q2]
        return r if d > 0 else r[:1]
    a0 = 1.0 / a0
    a1, a2, a3 = a1 * a0, a2 * a0, a3 * a0
    Q = (a1 * a1 - 3 * a2) * (1.0 / 9)
    R = (2 * a1 * a1 * a1 - 9 * a1 * a2 + 27 * a3) * (1.0 / 54)
    Qc = Q * Q * Q
    d = Qc - R * R
    if d > 0:
        theta = np.arccos(R / np.sqrt(Qc))
        t0, t1, t2 = -2 * np.sqrt(Q), theta / 3, a1 / 3
        return [t0 * np.cos(t1) - t2, t0 * np.cos(t1 + 2 * np.pi / 3) - t2, t0 * np.cos(t1 + 4 * np.pi / 3) - t2]
    if d == 0:
        if R >= 0:
            x0 = -2 * R ** (1.0 / 3) - a1 / 3
            x1 = R ** (1.0 / 3) - a1 / 3
        else:
            x0 = 2 * (-R) ** (1.0 / 3) - a1 / 3
            x1 = -((-R) ** (1.0 / 3)) - a1 / 3
        return [x0, x1, x1]
    d = np.sqrt(-d)
    e = (d + abs(R)) ** (1.0 / 3)
    if R > 0:
        e = -e
    return [(e + Q / e) - a1 / 3]


def run_7point(m1, m2):
    """FMEstimatorCallback::runKernel for 7 points (run7Point): normalise both point sets (centroid,
    mean distance sqrt 2), null space of the 7x9 system, det(lambda f1 + (1 - lambda) f2) = 0,
    de-normalise, scale to F[2,2] = 1.  Returns a list of 3x3 matrices (1..3)."""
    m1, m2 = np.asarray(m1, np.float64), np.asarray(m2, np.float64)
    c1, c2 = m1.mean(0), m2.mean(0)
    s1 = np.sqrt(((m1 - c1) ** 2).sum(1)).mean()
    s2 = np.sqrt(((m2 - c2) ** 2).sum(1)).mean()
    if s1 < np.finfo(np.float32).eps or s2 < np.finfo(np.float32).eps:
        return []
    s1, s2 = np.sqrt(2.0) / s1, np.sqrt(2.0) / s2
    a, b = (m1 - c1) * s1, (m2 - c2) * s2
    A = np.stack([b[:, 0] * a[:, 0], b[:, 0] * a[:, 1], b[:, 0], b[:, 1] * a[:, 0], b[:, 1] * a[:, 1], b[:, 1],
                  a[:, 0], a[:, 1], np.ones(7)], axis=1)
    _, _, Vt = np.linalg.svd(A, full_matrices=True)
    f1, f2 = Vt[7].copy(), Vt[8].copy()
    f1 -= f2
    A1, A2 = f1.reshape(3, 3), f2.reshape(3, 3)
    # det(l A1 + A2) as a cubic in l: exact interpolation through 4 points
    ls = np.array([-1.0, 0.0, 1.0, 2.0])
    dets = np.array([np.linalg.det(l * A1 + A2) for l in ls])
    c = np.linalg.solve(np.vander(ls, 4), dets)
    roots = solve_cubic(c)
    T1 = np.array([[s1, 0, -s1 * c1[0]], [0, s1, -s1 * c1[1]], [0, 0, 1]])
    T2 = np.array([[s2, 0, -s2 * c2[0]], [0, s2, -s2 * c2[1]], [0, 0, 1]])
    out = []
    for r in roots:
        lam, mu = r, 1.0
        s = f1[8] * r + f2[8]
        F = np.empty(9)
        if abs(s) > np.finfo(np.float64).eps:
            mu = 1.0 / s
            lam *= mu
            F[8] = 1.0
        else:
            F[8] = 0.0
        F[:8] = f1[:8] * lam + f2[:8] * mu
        F = T2.T @ F.reshape(3, 3) @ T1
        if abs(F[2, 2]) > np.finfo(np.float32).eps:
            F = F / F[2, 2]
        out.append(F)
    return out


def _cv2_ransac(p1, p2, model_points, kernel, error, thr, conf, max_iters, check_collinear, sample_pts=None):
    """RANSACPointSetRegistrator::run: rows of the sample table in order, every model of a sample scored,
    `goodCount > max(maxGoodCount, modelPoints - 1)` takes the lead (first wins ties), niters shrinks by
    RANSACUpdateNumIters.  Returns (model, mask bool[n], info)."""
    n = len(p1)
    s1, s2 = (p1, p2) if sample_pts is None else sample_pts
    rng_rows = cv2_sample_table(p1, p2, model_points, max_iters, check_collinear)
    niters, best_count, best, best_mask, best_iter = max(max_iters, 1), 0, None, None, -1
    t = np.float32(thr * thr)
    it = 0
    while it < niters and it < len(rng_rows):
        idx = rng_rows[it]
        for M in kernel(s1[idx], s2[idx]):
            mask = error(M, s1, s2) <= t
            good = int(mask.sum())
            if good > max(best_count, model_points - 1):
                best, best_mask, best_count, best_iter = M, mask, good, it
                niters = ransac_update_num_iters(conf, (n - good) / n, model_points, niters)
        it += 1
    return best, best_mask, dict(best_iter=best_iter, niters=niters, sample=rng_rows[best_iter] if best_iter >= 0 else None)


def ransac_fundamental_cv2(p1, p2, thr=3.0, conf=0.99, max_iters=1000):
    """cv2.findFundamentalMat(p1, p2, FM_RANSAC, thr, conf) (matcher.py:192, local_mapper.py:136): F [3,3] or None,
    mask bool [n], info.  With fewer than 15 points OpenCV silently runs its LMedS registrator instead."""
    p1, p2 = np.asarray(p1, np.float32), np.asarray(p2, np.float32)
    if len(p1) < 15:
        return lmeds_fundamental_cv2(p1, p2, conf, max_iters)
    return _cv2_ransac(p1, p2, 7, run_7point, fundamental_error, thr, conf, max_iters, True)


def lmeds_fundamental_cv2(p1, p2, conf=0.99, max_iters=1000, kernel=None):
    """LMeDSPointSetRegistrator::run for the fundamental matrix (what FM_RANSAC runs for 8..14 points): a fixed
    horizon for 45 % outliers, the model with the smallest median error (element n // 2 of the sorted errors;
    first wins ties), then inliers within sigma = 2.5 * 1.4826 * (1 + 5 / (n - 7)) * sqrt(median) (>= 0.001).
    F is None when fewer than 7 points pass that gate (cv2 returns an empty matrix, the mask stands)."""
    kernel = kernel or run_7point
    n = len(p1)
    niters = max(ransac_update_num_iters(conf, 0.45, 7, max_iters), 3)
    rows = cv2_sample_table(p1, p2, 7, niters, True, max_attempts=1000)
    best, min_med = None, np.inf
    for it, idx in enumerate(rows):
        for M in kernel(p1[idx], p2[idx]):
            med = float(np.sort(fundamental_error(M, p1, p2))[n // 2])
            if med < min_med:
                min_med, best = med, M
    if best is None:
        return None, np.zeros(n, bool), dict(niters=niters)
    sigma = max(2.5 * 1.4826 * (1 + 5.0 / (n - 7)) * np.sqrt(min_med), 0.001)
    mask = fundamental_error(best, p1, p2) <= np.float32(sigma * sigma)
    return (best if mask.sum() >= 7 else None), mask, dict(niters=niters, median=min_med)


def _poly_mul(a, b):
    """Product of two polynomials in (x, y, z) stored as dense [dx, dy, dz] coefficient arrays."""
    out = np.zeros(tuple(np.array(a.shape) + np.array(b.shape) - 1))
    for i, j, k in zip(*np.nonzero(a)):
        out[i:i + b.shape[0], j:j + b.shape[1], k:k + b.shape[2]] += a[i, j, k] * b
    return out


# the 20 monomials of degree <= 3 in (x, y, z): the ten that are led out first, then the basis of the quotient ring
_MONO = [(3, 0, 0), (2, 1, 0), (1, 2, 0), (0, 3, 0), (2, 0, 1), (1, 1, 1), (0, 2, 1), (1, 0, 2), (0, 1, 2), (0, 0, 3),
         (2, 0, 0), (1, 1, 0), (0, 2, 0), (1, 0, 1), (0, 1, 1), (0, 0, 2), (1, 0, 0), (0, 1, 0), (0, 0, 1), (0, 0, 0)]


def run_5point(q1, q2):
    """EMEstimatorCallback::runKernel (Nister's five-point problem): all real E with x2^T E x1 = 0 for the
    five correspondences, det E = 0 and 2 E E^T E - tr(E E^T) E = 0.  OpenCV eliminates to a degree-10
    polynomial in z and finds its roots with solvePoly; here the same ten cubic constraints are solved
    through the action matrix of x on the quotient ring (Stewenius), which has the same solution set.
    The ORDER of the solutions is OpenCV's only up to its own SVD basis and root finder; it matters
    only when two models of one sample tie on the inlier count at a lead change.  Returns unit-norm 3x3s."""
    q1, q2 = np.asarray(q1, np.float64), np.asarray(q2, np.float64)
    x1, y1, x2, y2 = q1[:, 0], q1[:, 1], q2[:, 0], q2[:, 1]
    Q = np.stack([x2 * x1, x2 * y1, x2, y2 * x1, y2 * y1, y2, x1, y1, np.ones(5)], axis=1)   # row-major E
    _, _, Vt = np.linalg.svd(Q, full_matrices=True)
    B = Vt[5:9]                                   # E = x B0 + y B1 + z B2 + B3
    # entries of E as polynomials of degree 1
    Ep = np.zeros((3, 3, 2, 2, 2))
    for r in range(3):
        for c in range(3):
            Ep[r, c, 1, 0, 0], Ep[r, c, 0, 1, 0], Ep[r, c, 0, 0, 1], Ep[r, c, 0, 0, 0] = B[:, 3 * r + c]
    mul = _poly_mul
    EEt = [[sum(mul(Ep[r, k], Ep[c, k]) for k in range(3)) for c in range(3)] for r in range(3)]
    tr = EEt[0][0] + EEt[1][1] + EEt[2][2]
    eqs = []
    for r in range(3):
        for c in range(3):
            eqs.append(2 * sum(mul(EEt[r][k], Ep[k, c]) for k in range(3)) - mul(tr, Ep[r, c]))
    det = (mul(mul(Ep[0, 0], Ep[1, 1]) - mul(Ep[0, 1], Ep[1, 0]), Ep[2, 2])
           - mul(mul(Ep[0, 0], Ep[1, 2]) - mul(Ep[0, 2], Ep[1, 0]), Ep[2, 1])
           + mul(mul(Ep[0, 1], Ep[1, 2]) - mul(Ep[0, 2], Ep[1, 1]), Ep[2, 0]))
    eqs.append(det)
    M = np.array([[e[m] for m in _MONO] for e in eqs])          # 10 x 20
    try:
        G = np.linalg.solve(M[:, :10], M[:, 10:])               # leading monomial i = -G[i] . basis
    except np.linalg.LinAlgError:
        return []
    # action matrices of x, y, z on the basis b = [x^2, xy, y^2, xz, yz, z^2, x, y, z, 1]: row i of A_v expresses
    # v * b[i] in the basis (leading monomials x^3 x^2y xy^2 y^3 x^2z xyz y^2z xz^2 yz^2 z^3 reduce through G)
    def action(lead_rows, basis_cols):
        A = np.zeros((10, 10))
        for i, r in enumerate(lead_rows):
            A[i] = -G[r]
        for i, c in zip(range(6, 10), basis_cols):
            A[i, c] = 1.0
        return A
    Ax, Ay, Az = action([0, 1, 2, 4, 5, 7], [0, 1, 3, 6]), action([1, 2, 3, 5, 6, 8], [1, 2, 4, 7]), \
        action([4, 5, 6, 7, 8, 9], [3, 4, 5, 8])

    def cons(c):
        E = (c @ B).reshape(3, 3)
        EEt = E @ E.T
        return np.r_[(2 * EEt @ E - np.trace(EEt) * E).ravel(), np.linalg.det(E), c @ c - 1.0]

    # candidates: eigenvectors of a few generic linear forms (A b(p) = form(p) b(p) at every solution p; a single
    # variable's action matrix has degenerate eigenvectors whenever two solutions share that coordinate)
    cands = []
    for wx, wy, wz in ((1.0, 0.0, 0.0), (0.3, 0.5, 0.8), (-0.7, 0.4, 0.6), (0.2, -0.9, 0.35)):
        w, V = np.linalg.eig(wx * Ax + wy * Ay + wz * Az)
        for i in range(10):
            v = V[:, i]
            if abs(v[9]) < 1e-14 * np.abs(v).max() or abs(w[i].imag) > 1e-3 * max(1.0, abs(w[i])):
                continue
            xyz = (v[6:9] / v[9]).real
            if np.all(np.isfinite(xyz)):
                c = np.r_[xyz, 1.0]
                cands.append(c / np.linalg.norm(c))
    # Gauss-Newton polish on the ten constraints (+ unit gauge) in null-space coordinates, then de-duplicate
    sols = []
    for c in cands:
        ok = False
        for _ in range(60):
            f = cons(c)
            J = np.empty((11, 4))
            for k in range(4):
                d = np.zeros(4)
                d[k] = 1e-7
                J[:, k] = (cons(c + d) - cons(c - d)) / 2e-7
            step = np.linalg.lstsq(J, -f, rcond=None)[0]
            c = c + step
            if np.linalg.norm(step) < 1e-14:
                ok = True
                break
        if not ok or np.linalg.norm(cons(c)) > 1e-11:
            continue
        if c[np.argmax(np.abs(c))] < 0:
            c = -c
        if all(np.linalg.norm(c - o) > 1e-7 for o in sols):
            sols.append(c)
    sols.sort(key=lambda c: c[2] / c[3] if c[3] != 0 else np.inf)     # ascending z, the variable OpenCV eliminates to
    return [(c @ B).reshape(3, 3) for c in sols]


def ransac_essential_cv2(p1, p2, K, prob=0.999, thr=1.0, max_iters=1000):
    """cv2.findEssentialMat(p1, p2, K, RANSAC, prob, thr) (utils.py:124-125, tracker.py:242): E [3,3] of unit
    Frobenius norm (cv2 returns norm sqrt 2... compare up to scale), mask bool [n], info."""
    K = np.asarray(K, np.float64)
    q1, q2 = normalise(p1, K), normalise(p2, K)
    t = thr / ((K[0, 0] + K[1, 1]) / 2)
    return _cv2_ransac(q1, q2, 5, run_5point, essential_error, t, prob, max_iters, False)
