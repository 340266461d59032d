"""Array-level front end for the callers either side of the hot path.

The reference's per-frame tracking step (`Tracker._track_frame`, tracker.py:214-254) is

    match -> filter_matches_by_geometric_distance -> filter_matches_by_distance
          -> cv2.findEssentialMat(RANSAC, 0.999, 1.0) -> cv2.recoverPose

written with Python lists of cv2.DMatch / cv2.KeyPoint (about 4 ms of object construction per frame).
`track_pair` is the same sequence on plain arrays: the kNN-2 search, the RANSAC and the pose recovery run in
libvslam_b200.so, the two gates are the reference's arithmetic vectorised (exactly the same decisions: the
comparisons are made on the same float64 values).  No cv2 object is created on the way.

`BFMatcher` / `install_local_mapper_adaptor` route LocalMapper's private matcher (local_mapper.py:116-149:
its own `cv2.BFMatcher.knnMatch`, a 0.8 ratio test and `cv2.findFundamentalMat(FM_RANSAC, 3.0)`) through the
same kernels without editing local_mapper.py.
"""
import math

import numpy as np

from . import _native


# --------------------------------------------------------------------------- N1: tracking step
def match_arrays(descriptors1, descriptors2, ratio_threshold=0.75, ratio_test=True):
    """matcher.py:44-83 on arrays -> (query_idx int32[n], train_idx int32[n], distance float32[n]),
    ascending query index (the order of the reference's match list)."""
    e = np.zeros(0, np.int32)
    if descriptors1 is None or descriptors2 is None or len(descriptors1) == 0 or len(descriptors2) == 0:
        return e, e.copy(), np.zeros(0, np.float32)
    idx, dist = _native.bf_knn2_host(np.asarray(descriptors1).astype(np.uint8, copy=False),
                                     np.asarray(descriptors2).astype(np.uint8, copy=False))
    d1 = dist[:, 0].astype(np.float32).astype(np.float64)
    d2 = dist[:, 1].astype(np.float32).astype(np.float64)
    two = idx[:, 1] >= 0
    one = (idx[:, 0] >= 0) & ~two
    keep = ((two & (d1 < ratio_threshold * d2)) | one) if ratio_test else (two | one)
    q = np.nonzero(keep)[0].astype(np.int32)
    return q, idx[q, 0].astype(np.int32), dist[q, 0].astype(np.float32)


def gate_geometric_distance(pts1, pts2, q, t, threshold_percent, image_shape):
    """matcher.py:109-142: keep matches whose displacement is <= ((W + H) / 2) * pct.  pts are float32 [n, 2]
    (cv2.KeyPoint.pt values); the differences and the hypot are taken in float64 like the reference's
    Python floats.  Returns the boolean keep mask over the matches."""
    height, width = image_shape[:2]
    max_dist = ((width + height) / 2.0) * threshold_percent
    p1 = np.asarray(pts1, np.float32)[q].astype(np.float64)
    p2 = np.asarray(pts2, np.float32)[t].astype(np.float64)
    dx, dy = p2[:, 0] - p1[:, 0], p2[:, 1] - p1[:, 1]
    d = np.hypot(dx, dy)
    keep = d <= max_dist
    # math.hypot (the reference) and C hypot may differ in the last ulp: re-decide the borderline ones exactly
    for k in np.nonzero(np.abs(d - max_dist) <= 1e-9 * max(max_dist, 1.0))[0]:
        keep[k] = math.hypot(dx[k], dy[k]) <= max_dist
    return keep


def gate_descriptor_distance(dist, distance_threshold=None):
    """matcher.py:144-169: stable sort by distance, threshold = 2 x median when not given, keep `<`.
    Returns (order, n_keep): indices into the match arrays in the reference's output order."""
    dist = np.asarray(dist, np.float32)
    if len(dist) == 0:
        return np.zeros(0, np.int64), 0
    order = np.argsort(dist, kind="stable")
    if distance_threshold is None:
        distance_threshold = np.median(dist.astype(np.float64)) * 2.0
    n_keep = int(np.count_nonzero(dist[order].astype(np.float64) < distance_threshold))
    return order, n_keep


def track_pair(pts1, descriptors1, pts2, descriptors2, camera_matrix, image_shape, ratio_threshold=0.75,
               threshold_percent=0.02, prob=0.999, threshold=1.0):
    """tracker.py:214-254 on arrays.

    pts*: float32 [n, 2] keypoint coordinates (level-0 pixels), descriptors*: uint8 [n, 32].
    Returns a dict: success, query_idx / train_idx / distance (the gated matches in the reference's order,
    i.e. ascending descriptor distance, stable), E, ransac_mask (uint8 [m], findEssentialMat's), inlier_mask
    (uint8 [m], after recoverPose's cheirality check), R, t, T (4x4).
    `success` is False, with the matches still reported, where the reference returns (False, None, None).
    """
    from .utils import calculate_essential_matrix, recover_pose
    out = {"success": False, "E": None, "R": None, "t": None, "T": None, "ransac_mask": None, "inlier_mask": None}
    q, t, d = match_arrays(descriptors1, descriptors2, ratio_threshold)
    if len(q):
        keep = gate_geometric_distance(pts1, pts2, q, t, threshold_percent, image_shape)
        q, t, d = q[keep], t[keep], d[keep]
    order, n_keep = gate_descriptor_distance(d)
    sel = order[:n_keep]
    q, t, d = q[sel], t[sel], d[sel]
    out.update(query_idx=q, train_idx=t, distance=d)
    if len(q) < 8:                                             # tracker.py:233-234
        return out
    p1 = np.ascontiguousarray(np.asarray(pts1, np.float32)[q])
    p2 = np.ascontiguousarray(np.asarray(pts2, np.float32)[t])
    E, mask = calculate_essential_matrix(p1, p2, camera_matrix, prob=prob, threshold=threshold)
    if E is None or E.shape != (3, 3):                         # tracker.py:244-245
        return out
    out["ransac_mask"] = mask.ravel().copy()
    _, R, tv, mask = recover_pose(E, p1, p2, camera_matrix, mask=mask)
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = tv.reshape(3)
    out.update(success=True, E=E, R=R, t=tv, T=T, inlier_mask=mask.ravel())
    return out


# --------------------------------------------------------------------------- full front end, batched
def front_end_full(handle, frames, ratio=0.75, threshold=3.0, confidence=0.99, max_iters=1000, seed=0):
    """BASELINE config 3 ("full front-end"): extract every frame, match frame i with i+1, and filter every
    match list with the fundamental-matrix RANSAC (matcher.py:171-200) -- all on the device, one host
    round trip.  `handle` is an `_native.OrbHandle`, `frames` uint8 [B, H, W].

    Returns numpy arrays: kp (structured, [B, cap]), desc uint8 [B, cap, 32], count int32 [B],
    match int32 [B-1, cap, 3], match_count int32 [B-1], F float64 [B-1, 3, 3], mask uint8 [B-1, cap],
    n_inliers int32 [B-1] (-1: fewer than 8 matches, mask rows all 1 as in matcher.py:184-185).
    """
    import ctypes as C
    import torch
    _native.require_gpu()
    L = _native.lib()
    frames = np.ascontiguousarray(frames, np.uint8)
    B, h, w = frames.shape
    cap = handle.kp_capacity
    dev = torch.device("cuda")
    d_frames = torch.from_numpy(frames).to(dev)
    kp = torch.empty((B, cap, 6), dtype=torch.float32, device=dev)
    desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    count = torch.zeros(B, dtype=torch.int32, device=dev)
    pairs = max(B - 1, 0)
    match = torch.empty((max(pairs, 1), cap, 3), dtype=torch.int32, device=dev)
    mcount = torch.zeros(max(pairs, 1), dtype=torch.int32, device=dev)
    F = torch.zeros((max(pairs, 1), 9), dtype=torch.float64, device=dev)
    mask = torch.zeros((max(pairs, 1), cap), dtype=torch.uint8, device=dev)
    ninl = torch.zeros(max(pairs, 1), dtype=torch.int32, device=dev)
    ws = torch.empty(max(int(L.vslam_ransac_fundamental_batched_workspace(pairs, cap)), 16), dtype=torch.uint8, device=dev)
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _native.check(L.vslam_frontend_run(handle.handle, d_frames.data_ptr(), B, w, None, None, float(ratio),
                                       kp.data_ptr(), desc.data_ptr(), count.data_ptr(), match.data_ptr(),
                                       mcount.data_ptr(), s))
    if pairs:
        _native.check(L.vslam_ransac_fundamental_batched(kp.data_ptr(), kp.data_ptr() + cap * 24, cap, match.data_ptr(),
                                                         mcount.data_ptr(), cap, pairs, float(threshold),
                                                         float(confidence), int(max_iters), C.c_uint64(seed),
                                                         ws.data_ptr(), F.data_ptr(), mask.data_ptr(), ninl.data_ptr(), s))
    torch.cuda.synchronize()
    _native.check(L.vslam_orb_last_status(handle.handle))       # retain-best tie overflow -> VSLAM_E_CAPACITY, loudly
    kp_np = kp.cpu().numpy().view(_native.KP_DTYPE).reshape(B, cap)
    return {"kp": kp_np, "desc": desc.cpu().numpy(), "count": count.cpu().numpy().clip(max=cap),
            "match": match[:pairs].cpu().numpy(), "match_count": mcount[:pairs].cpu().numpy(),
            "F": F[:pairs].cpu().numpy().reshape(pairs, 3, 3), "mask": mask[:pairs].cpu().numpy(),
            "n_inliers": ninl[:pairs].cpu().numpy()}


def track_batched(handle, frames, camera_matrix, ratio=0.75, threshold_percent=0.02, prob=0.999, threshold=1.0,
                  max_iters=1000, seed=None):
    """The reference's tracking step (tracker.py:214-254) for every consecutive frame pair of `frames`
    (uint8 [B, H, W]) in two launches after the extraction: kNN-2 + ratio (vslam_frontend_run), then
    vslam_track_batched (both gates, 5-point RANSAC, recoverPose -- one CTA per pair).  Pair p equals
    `track_pair` on frames p, p+1; a pair whose displacement gate had a decision within 1e-9 of the limit is
    re-run through `track_pair` (math.hypot on the host), which has not happened on any test input.

    Returns numpy arrays: kp, desc, count as `front_end_full`; track_match int32 [B-1, cap, 3] /
    track_count int32 [B-1] (gated matches, ascending distance, stable); E, R float64 [B-1, 3, 3];
    t float64 [B-1, 3, 1]; T float64 [B-1, 4, 4]; ransac_mask, inlier_mask uint8 [B-1, cap];
    success bool [B-1]; status int32 [B-1, 4] (include/vslam_b200.h)."""
    import ctypes as C
    import torch
    from .utils import DEFAULT_SEED
    _native.require_gpu()
    L = _native.lib()
    if seed is None:
        seed = DEFAULT_SEED
    frames = np.ascontiguousarray(frames, np.uint8)
    B, h, w = frames.shape
    cap = handle.kp_capacity
    dev = torch.device("cuda")
    d_frames = torch.from_numpy(frames).to(dev)
    kp = torch.empty((B, cap, 6), dtype=torch.float32, device=dev)
    desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    count = torch.zeros(B, dtype=torch.int32, device=dev)
    pairs = max(B - 1, 0)
    P = max(pairs, 1)
    match = torch.empty((P, cap, 3), dtype=torch.int32, device=dev)
    mcount = torch.zeros(P, dtype=torch.int32, device=dev)
    tmatch = torch.zeros((P, cap, 3), dtype=torch.int32, device=dev)
    tcount = torch.zeros(P, dtype=torch.int32, device=dev)
    E = torch.zeros((P, 9), dtype=torch.float64, device=dev)
    R = torch.zeros((P, 9), dtype=torch.float64, device=dev)
    t = torch.zeros((P, 3), dtype=torch.float64, device=dev)
    rmask = torch.zeros((P, cap), dtype=torch.uint8, device=dev)
    pmask = torch.zeros((P, cap), dtype=torch.uint8, device=dev)
    status = torch.zeros((P, 4), dtype=torch.int32, device=dev)
    ws = torch.empty(max(int(L.vslam_track_batched_workspace(pairs, cap)), 16), dtype=torch.uint8, device=dev)
    K = np.ascontiguousarray(camera_matrix, np.float64).reshape(3, 3)
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _native.check(L.vslam_frontend_run(handle.handle, d_frames.data_ptr(), B, w, None, None, float(ratio),
                                       kp.data_ptr(), desc.data_ptr(), count.data_ptr(), match.data_ptr(),
                                       mcount.data_ptr(), s))
    if pairs:
        _native.check(L.vslam_track_batched(kp.data_ptr(), kp.data_ptr() + cap * 24, cap, match.data_ptr(),
                                            mcount.data_ptr(), cap, pairs, _native.ptr(K), w, h,
                                            float(threshold_percent), float(prob), float(threshold), int(max_iters),
                                            C.c_uint64(seed), ws.data_ptr(), tmatch.data_ptr(), tcount.data_ptr(),
                                            E.data_ptr(), rmask.data_ptr(), pmask.data_ptr(), R.data_ptr(),
                                            t.data_ptr(), status.data_ptr(), s))
    torch.cuda.synchronize()
    _native.check(L.vslam_orb_last_status(handle.handle))       # retain-best tie overflow -> VSLAM_E_CAPACITY, loudly
    kp_np = kp.cpu().numpy().view(_native.KP_DTYPE).reshape(B, cap)
    out = {"kp": kp_np, "desc": desc.cpu().numpy(), "count": count.cpu().numpy().clip(max=cap),
           "track_match": tmatch[:pairs].cpu().numpy(), "track_count": tcount[:pairs].cpu().numpy(),
           "E": E[:pairs].cpu().numpy().reshape(pairs, 3, 3), "R": R[:pairs].cpu().numpy().reshape(pairs, 3, 3),
           "t": t[:pairs].cpu().numpy().reshape(pairs, 3, 1), "ransac_mask": rmask[:pairs].cpu().numpy(),
           "inlier_mask": pmask[:pairs].cpu().numpy(), "status": status[:pairs].cpu().numpy()}
    out["success"] = out["status"][:, 0] == 1 if pairs else np.zeros(0, bool)
    T = np.tile(np.eye(4), (pairs, 1, 1))
    if pairs:
        T[:, :3, :3] = out["R"]
        T[:, :3, 3] = out["t"][:, :, 0]
    out["T"] = T
    # borderline displacement decisions: the host's math.hypot decides (matcher.py:138)
    for p in np.nonzero(out["status"][:, 3] > 0)[0] if pairs else []:
        n0, n1 = int(out["count"][p]), int(out["count"][p + 1])
        pts0 = np.stack([kp_np[p]["x"][:n0], kp_np[p]["y"][:n0]], 1)
        pts1 = np.stack([kp_np[p + 1]["x"][:n1], kp_np[p + 1]["y"][:n1]], 1)
        r = track_pair(pts0, out["desc"][p, :n0], pts1, out["desc"][p + 1, :n1], K, (h, w), ratio, threshold_percent,
                       prob, threshold)
        m = len(r["query_idx"])
        out["track_count"][p] = m
        out["track_match"][p, :m, 0] = r["query_idx"]
        out["track_match"][p, :m, 1] = r["train_idx"]
        out["track_match"][p, :m, 2] = r["distance"].astype(np.int32)
        out["success"][p] = r["success"]
        if r["success"]:
            out["E"][p], out["R"][p], out["t"][p], out["T"][p] = r["E"], r["R"], r["t"], r["T"]
            out["ransac_mask"][p, :m] = r["ransac_mask"]
            out["inlier_mask"][p, :m] = r["inlier_mask"]
    return out


def track_clip(handle, frames, camera_matrix, ratio=0.75, threshold_percent=0.02, prob=0.999, threshold=1.0,
               max_iters=1000, seed=None):
    """Tracker.process_frame's loop (tracker.py:73-146) over a clip, one library call: `frames` uint8 [B, H, W]
    go up once, and only the relative poses come back -> (T float64 [B-1, 4, 4], success bool [B-1],
    status int32 [B-1, 4], count int32 [B]).  Pair p equals `track_batched` / `track_pair` on frames p, p+1
    (a pair whose status[3] is non-zero had a displacement decision within 1e-9 of the limit and should be
    re-run through `track_pair`)."""
    import ctypes as C
    from .utils import DEFAULT_SEED
    _native.require_gpu()
    frames = np.ascontiguousarray(frames, np.uint8)
    B = frames.shape[0]
    K = np.ascontiguousarray(camera_matrix, np.float64).reshape(3, 3)
    T = np.zeros((B - 1, 4, 4), np.float64)
    status = np.zeros((B - 1, 4), np.int32)
    count = np.zeros(B, np.int32)
    _native.check(_native.lib().vslam_track_clip_host(handle.handle, _native.ptr(frames), B, float(ratio), _native.ptr(K),
                                                      float(threshold_percent), float(prob), float(threshold),
                                                      int(max_iters), C.c_uint64(DEFAULT_SEED if seed is None else seed),
                                                      _native.ptr(T), _native.ptr(status), _native.ptr(count)))
    return T, status[:, 0] == 1, status, count


# --------------------------------------------------------------------------- N2: LocalMapper's matcher
class BFMatcher:
    """Stand-in for `cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False)` (local_mapper.py:116): `knnMatch`
    with k = 1 or 2 and `match` on the B200 kernels, returning the same tuples of cv2.DMatch."""

    def __init__(self, normType=None, crossCheck=False):
        import cv2
        if normType not in (None, cv2.NORM_HAMMING):
            raise NotImplementedError("only NORM_HAMMING runs on the B200 path")
        if crossCheck:
            raise NotImplementedError("crossCheck=True is not used by the reference")

    @staticmethod
    def _search(query, train):
        if query is None or train is None or len(query) == 0:
            return np.zeros((0, 2), np.int32), np.zeros((0, 2), np.int32)
        if len(train) == 0:
            n = len(query)
            return np.full((n, 2), -1, np.int32), np.full((n, 2), -1, np.int32)
        return _native.bf_knn2_host(np.asarray(query).astype(np.uint8, copy=False),
                                    np.asarray(train).astype(np.uint8, copy=False))

    def knnMatch(self, queryDescriptors, trainDescriptors, k=2, mask=None, compactResult=False):
        import cv2
        if mask is not None or k not in (1, 2):
            raise NotImplementedError("knnMatch: k in (1, 2) without a mask")
        idx, dist = self._search(queryDescriptors, trainDescriptors)
        out = []
        for r in range(len(idx)):
            row = tuple(cv2.DMatch(r, int(idx[r, j]), 0, float(dist[r, j])) for j in range(k) if idx[r, j] >= 0)
            out.append(row)
        return tuple(out)

    def match(self, queryDescriptors, trainDescriptors, mask=None):
        return tuple(m[0] for m in self.knnMatch(queryDescriptors, trainDescriptors, k=1, mask=mask) if m)


class _Cv2Proxy:
    """cv2 as seen by an adapted module: BFMatcher and findFundamentalMat(FM_RANSAC) run on the B200
    kernels, every other attribute is cv2's own."""

    def __init__(self, cv2):
        self._cv2 = cv2

    def __getattr__(self, name):
        return getattr(self._cv2, name)

    def BFMatcher(self, normType=None, crossCheck=False):
        return BFMatcher(normType, crossCheck)

    def findFundamentalMat(self, points1, points2, method=None, ransacReprojThreshold=3.0, confidence=0.99,
                           maxIters=1000, mask=None):
        cv2 = self._cv2
        if method is None:
            method = cv2.FM_RANSAC
        n = len(points1)
        if method != cv2.FM_RANSAC or n < 8:
            # no CPU fallback in this package: the reference's callers use FM_RANSAC on >= 8 matches only
            # (local_mapper.py:136 after its own len check, matcher.py:184-192)
            raise NotImplementedError("findFundamentalMat on the B200 kernels: FM_RANSAC with at least 8 points only")
        from .utils import find_fundamental_ransac
        return find_fundamental_ransac(np.asarray(points1, np.float32).reshape(-1, 2),
                                       np.asarray(points2, np.float32).reshape(-1, 2),
                                       ransacReprojThreshold, confidence, maxIters)


    def findEssentialMat(self, points1, points2, cameraMatrix=None, method=None, prob=0.999, threshold=1.0,
                         maxIters=1000, mask=None):
        """cv2.findEssentialMat(p1, p2, K, RANSAC, prob, threshold) as Tracker._track_from_last_frame calls it
        (tracker.py:242) -> (E float64 [3,3] or None, mask uint8 [n,1])."""
        cv2 = self._cv2
        if method is None:
            method = cv2.RANSAC
        if method != cv2.RANSAC or cameraMatrix is None or len(points1) < 5:
            raise NotImplementedError("findEssentialMat on the B200 kernels: RANSAC with a camera matrix, >= 5 points")
        from .utils import calculate_essential_matrix
        return calculate_essential_matrix(points1, points2, cameraMatrix, prob=prob, threshold=threshold,
                                          max_iters=maxIters)

    def recoverPose(self, E, points1, points2, cameraMatrix=None, mask=None, **kw):
        """cv2.recoverPose(E, p1, p2, K, mask=mask) (tracker.py:249) -> (n_inliers, R, t, mask)."""
        if cameraMatrix is None or kw:
            raise NotImplementedError("recoverPose on the B200 kernels: the (E, points1, points2, cameraMatrix, mask) form")
        from .utils import recover_pose
        return recover_pose(E, points1, points2, cameraMatrix, mask)


def install_tracker_adaptor(module):
    """Point `module.cv2` (the reference's `orbslam2.tracker`) at the proxy: Tracker._track_from_last_frame calls
    cv2.findEssentialMat / cv2.recoverPose directly (tracker.py:242,249), so swapping the extractor / matcher /
    initializer classes alone leaves those two on the CPU.  Drawing calls stay cv2's.  Undo: module.cv2 = proxy._cv2."""
    import cv2
    _native.require_gpu()
    proxy = _Cv2Proxy(cv2)
    module.cv2 = proxy
    return proxy


def install_local_mapper_adaptor(module):
    """Point `module.cv2` (e.g. the reference's `orbslam2.local_mapper`) at the proxy above, so that its
    private matcher and fundamental-matrix RANSAC use the B200 kernels, and -- when the module defines
    `LocalMapper` -- replace `LocalMapper._cull_map_points` by the device version.  Returns the proxy; undo
    with `module.cv2 = proxy._cv2` (and `module.LocalMapper._cull_map_points = proxy._cull`)."""
    import cv2
    _native.require_gpu()
    proxy = _Cv2Proxy(cv2)
    module.cv2 = proxy
    cls = getattr(module, "LocalMapper", None)
    if cls is not None and hasattr(cls, "_cull_map_points"):
        proxy._cull = cls._cull_map_points
        cls._cull_map_points = _cull_map_points_method
    return proxy


# --------------------------------------------------------------------------- N2: LocalMapper's map-point culling
def reprojection_cull(points, obs_offset, obs_keyframe, obs_xy, projections, min_observations=2, max_error=5.0):
    """The per-point test of LocalMapper._cull_map_points (local_mapper.py:205-237) on flat arrays, on the GPU:
    points [n, 3]; CSR observation lists obs_offset [n + 1], obs_keyframe [m], obs_xy [m, 2] (kp.pt);
    projections [n_kf, 3, 4] = K [R | t] per keyframe.  -> (valid bool [n], err float64 [n])."""
    _native.require_gpu()
    pts = np.ascontiguousarray(np.asarray(points, np.float64).reshape(-1, 3))
    off = np.ascontiguousarray(obs_offset, np.int32)
    kf = np.ascontiguousarray(obs_keyframe, np.int32)
    xy = np.ascontiguousarray(np.asarray(obs_xy, np.float32).reshape(-1, 2))
    P = np.ascontiguousarray(np.asarray(projections, np.float64).reshape(-1, 12))
    n = len(pts)
    if len(off) != n + 1 or len(kf) != len(xy) or (n and int(off[-1]) != len(kf)):
        raise ValueError("observation lists do not match the points")
    valid, err = np.zeros(n, np.uint8), np.zeros(n, np.float64)
    _native.check(_native.lib().vslam_cull_map_points_host(_native.ptr(pts), _native.ptr(off), _native.ptr(kf),
                                                           _native.ptr(xy), _native.ptr(P), n, len(P),
                                                           int(min_observations), float(max_error),
                                                           _native.ptr(valid), _native.ptr(err)))
    return valid == 1, err


def cull_map_points(map_points, keyframes, camera_matrix, min_observations=2, max_error=5.0):
    """LocalMapper._cull_map_points' first half on the reference's own structures: `map_points` is the list of
    dicts with 'position' and 'observed_keyframes' {keyframe index: keypoint index}, `keyframes` the list of
    dicts with 'keypoints' (cv2.KeyPoint) and 'pose' (4x4).  Returns the surviving dicts, in order."""
    from .utils import compute_projection_matrix
    if not map_points:
        return []
    P = np.zeros((len(keyframes), 3, 4))
    for k, kf in enumerate(keyframes):
        P[k] = compute_projection_matrix(kf['pose'][:3, :3], kf['pose'][:3, 3], camera_matrix)
    off, okf, oxy = [0], [], []
    for mp in map_points:
        for kf_id, kp_id in mp.get('observed_keyframes', {}).items():
            okf.append(kf_id)
            oxy.append(keyframes[kf_id]['keypoints'][kp_id].pt)      # IndexError like the reference's lookup
        off.append(len(okf))
    pos = np.array([np.asarray(mp['position'], np.float64).reshape(3) for mp in map_points])
    valid, _ = reprojection_cull(pos, off, okf, np.asarray(oxy, np.float32).reshape(-1, 2), P, min_observations, max_error)
    return [mp for mp, v in zip(map_points, valid) if v]


def _cull_map_points_method(self):
    """Drop-in body for LocalMapper._cull_map_points (local_mapper.py:205-245): the per-point test on the GPU,
    then the per-keyframe id lists rebuilt like local_mapper.py:240-245 (same order, one pass)."""
    self.map_points = cull_map_points(self.map_points, self.keyframes, self.camera_matrix, self.min_observations)
    lists = {kf['id']: [] for kf in self.keyframes}
    for mp in self.map_points:
        for kf_id in mp.get('observed_keyframes', {}):
            if kf_id in lists:
                lists[kf_id].append(mp['id'])
    for kf in self.keyframes:
        kf['map_points'] = lists[kf['id']]
