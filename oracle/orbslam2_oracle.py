"""CPU oracle of the ORB-SLAM2-style keypoint distribution.

TEST INFRASTRUCTURE ONLY (imported by tests/ and nothing else).

PARITY UNPINNED: the reference stores `ini_threshold` / `min_threshold` (src/orbslam2/extractor.py:19-35) but
never implements what they mean in ORB-SLAM2 -- it calls cv2.ORB (extractor.py:38-48,65).  There is therefore
no reference output, test or golden vector for this mode.  This file restates the published algorithm
(Mur-Artal & Tardos, ORB-SLAM2, 2017: ORBextractor::ComputeKeyPointsOctTree and ::DistributeOctTree) the way
the original is written -- a list of nodes that are split one by one -- and the CUDA kernel
(csrc/orb_octree.cu, which works from sorted path keys instead) must reproduce it exactly.

Stated deviations from ORB-SLAM2, shared with the kernel:
  * pyramid, blur, orientation and rBRIEF are those of the cv2.ORB path (oracle/orb_oracle.c), so the edge is
    cv2.ORB's 31 px instead of 19 and FAST's non-max suppression runs over the whole level, not per cell;
  * a round always splits the fullest nodes first, ties in quadtree (Z) order; ORB-SLAM2 orders only its last
    round and breaks ties by pointer value (not reproducible).  Rounds that cannot reach the quota split every
    node, so the order does not matter there and the node sets are identical;
  * more than 16384 corners on one level: the weakest are dropped first (threshold from the score histogram).
"""
import numpy as np

from . import oracle as O

EDGE = 31
CELL = 30
MAXM = 16384


def cell_filter(xs, ys, ss, w, h, ini_thr):
    """ComputeKeyPointsOctTree's per-cell threshold: a 30-px cell that holds a corner with score >= ini_thr
    keeps only those; a cell without one keeps everything FAST found at the minimum threshold.  Returns the
    boolean keep mask."""
    minb = EDGE - 3
    width, height = w - 2 * minb, h - 2 * minb
    ncols, nrows = max(width // CELL, 1), max(height // CELL, 1)
    wcell, hcell = -(-width // ncols), -(-height // nrows)
    tcap = 0
    if len(xs) > MAXM:
        hist = np.bincount(ss, minlength=256)
        acc = 0
        for t in range(255, 0, -1):
            acc += int(hist[t])
            if acc > MAXM:
                tcap = t + 1
                break
    cj, ci = (xs - minb - 3) // wcell, (ys - minb - 3) // hcell
    strong = ss >= ini_thr
    cells = set(zip(ci[strong].tolist(), cj[strong].tolist()))
    has_ini = np.array([(a, b) in cells for a, b in zip(ci.tolist(), cj.tolist())], bool) if len(xs) else strong
    skipped = (cj * wcell >= width - 6) | (ci * hcell >= height - 3)
    return ~skipped & (strong | ~has_ini) & (ss >= tcap)


def distribute_octree(rx, ry, score, width, height, n_target):
    """DistributeOctTree on coordinates relative to the region's corner.  Returns the indices of the kept
    corners in quadtree order."""
    if len(rx) == 0 or n_target <= 0:
        return []
    nini = min(max(int(np.floor(np.float32(width) / np.float32(height) + np.float32(0.5))), 1), 255)
    hx = np.float32(width) / np.float32(nini)
    nodes = {}
    for i in range(len(rx)):
        r = min(int(np.float32(rx[i]) / hx), nini - 1)
        nodes.setdefault((r,), dict(b=[int(hx * np.float32(r)), int(hx * np.float32(r + 1)), 0, height], pts=[]))
        nodes[(r,)]["pts"].append(i)
    done, first = False, True
    while not done:
        prev = len(nodes)
        if not first and prev >= n_target:
            break
        order = sorted((p for p in nodes if len(nodes[p]["pts"]) > 1), key=lambda p: (-len(nodes[p]["pts"]), p))
        if not order:
            break
        for p in order:
            nd = nodes.pop(p)
            lox, hix, loy, hiy = nd["b"]
            mx, my = lox + (hix - lox + 1) // 2, loy + (hiy - loy + 1) // 2
            for i in nd["pts"]:
                cx, cy = int(rx[i] >= mx), int(ry[i] >= my)
                q = p + (cx + 2 * cy,)
                if q not in nodes:
                    nodes[q] = dict(b=[mx if cx else lox, hix if cx else mx, my if cy else loy, hiy if cy else my], pts=[])
                nodes[q]["pts"].append(i)
            if not first and len(nodes) >= n_target:     # the first round splits every root node (ORB-SLAM2's loop
                done = True                              # only orders and cuts short the rounds after it)
                break
        first = False
        if len(nodes) == prev:
            break
    out = []
    for p in sorted(nodes):
        # corners of a node in quadtree order (deeper path), first maximum wins
        pts = sorted(nodes[p]["pts"], key=lambda i: _path(rx[i], ry[i], nodes[p]["b"]))
        best = pts[0]
        for i in pts[1:]:
            if score[i] > score[best]:
                best = i
        out.append(best)
    return out


def _path(x, y, b):
    lox, hix, loy, hiy = b
    p = []
    for _ in range(14):
        mx, my = lox + (hix - lox + 1) // 2, loy + (hiy - loy + 1) // 2
        cx, cy = int(x >= mx), int(y >= my)
        lox, hix = (mx, hix) if cx else (lox, mx)
        loy, hiy = (my, hiy) if cy else (loy, my)
        p.append(cx + 2 * cy)
    return tuple(p)


def detect_and_compute(img, nfeatures=2000, scale_factor=1.2, nlevels=8, ini_thr=20, min_thr=7):
    """The extractor in ORB-SLAM2 distribution mode -> dict like oracle.orb_detect_and_compute."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    lw, lh, ls = O.level_geometry(w, h, scale_factor, nlevels)
    quota = O.level_quotas(nfeatures, scale_factor, nlevels)
    X, Y, S, A, R, OC, LX, LY = [], [], [], [], [], [], [], []
    counts = np.zeros(nlevels, np.int32)
    for l in range(nlevels):
        xs, ys, ss = O.level_fast_candidates(img, l, scale_factor, nlevels, min_thr)
        if len(xs) == 0:
            continue
        keep = cell_filter(xs, ys, ss, int(lw[l]), int(lh[l]), ini_thr)
        xs, ys, ss = xs[keep], ys[keep], ss[keep]
        minb = EDGE - 3
        sel = distribute_octree(xs - minb, ys - minb, ss, int(lw[l]) - 2 * minb, int(lh[l]) - 2 * minb, int(quota[l]))
        raw = O.pyramid_level(img, l, scale_factor, nlevels) if l > 0 else img
        sc = np.float32(ls[l])
        for i in sel:
            X.append(np.float32(xs[i]) * sc); Y.append(np.float32(ys[i]) * sc)
            S.append(np.float32(31.0) * sc); A.append(np.float32(O.ic_angle(raw, xs[i], ys[i])))
            R.append(np.float32(ss[i])); OC.append(l); LX.append(xs[i]); LY.append(ys[i])
        counts[l] = len(sel)
    x, y = np.array(X, np.float32), np.array(Y, np.float32)
    ang, octv = np.array(A, np.float32), np.array(OC, np.int32)
    if len(x):
        desc, keepi = O.orb_compute(img, x, y, ang, octv, scale_factor)
        assert len(keepi) == len(x)
    else:
        desc = np.zeros((0, 32), np.uint8)
    return dict(x=x, y=y, size=np.array(S, np.float32), angle=ang, response=np.array(R, np.float32), octave=octv,
                lx=np.array(LX, np.int32), ly=np.array(LY, np.int32), desc=desc, level_counts=counts)
