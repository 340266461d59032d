"""bench.py --latency: what one call costs, for the calls the reference's callers make per frame.

BASELINE.json config 2 is worded "single-frame ORB extract + consecutive-frame match": the throughput line of
bench.py batches 512 frames per launch, the reference's Tracker handles ONE frame per call
(src/orbslam2/tracker.py:87,214).  This mode reports p50 / p99 over >= 200 calls at C2 (640x480, 2000 features,
1 GPU) of
  (i)   the C ABI, one frame per call: pinned frame in -> vslam_frontend_run(batch 1, previous frame's descriptors
        on the device) -> keypoints, descriptors and the match list back in pinned memory -> synchronise;
        `kernel_us` is the same call timed by CUDA events around the kernels only;
  (ii)  the drop-in object API: ORBExtractor.detect_and_compute + DescriptorMatcher.match (cv2.KeyPoint / DMatch
        objects built, as the reference's callers receive them);
  (iii) ORBExtractor.distribute_keypoints (K8: the path Tracker.process_frame really takes, extractor.py:85-144);
each beside the time of the reference's cv2 calls for the same work on this box's host cores.
One JSON line on stdout.
"""
import ctypes as C
import json
import os
import time

import numpy as np


def _pcts(us):
    a = np.sort(np.asarray(us, np.float64))
    return {"p50_us": float(a[len(a) // 2]), "p99_us": float(a[min(len(a) - 1, int(len(a) * 0.99))]),
            "mean_us": float(a.mean()), "calls": int(len(a))}


def run_latency(args):
    import torch
    import bench as BN
    from orbslam2 import _native
    from orbslam2.extractor import ORBExtractor
    from orbslam2.matcher import DescriptorMatcher
    from orbslam2.synth import SyntheticVideo

    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L = _native.lib()
    w, h, nf, _ = BN.WORKLOADS["c2"]
    n_calls = max(200, args.steps)
    video = SyntheticVideo(w, h, 0)
    frames = video.frames(0, 32)
    out = {"mode": "latency", "workload": BN.workload_name("c2"), "n_gpus": 1, "calls": n_calls}

    # ---- (i) C ABI, one frame per call -------------------------------------------------------------------------
    orb = _native.OrbHandle(w, h, n_features=nf, scale_factor=BN.SCALE, n_levels=BN.N_LEVELS,
                            fast_threshold=BN.FAST_THR, max_batch=1)
    cap = orb.kp_capacity
    pin_in = torch.from_numpy(frames).pin_memory()
    d_frame = torch.empty((1, h, w), dtype=torch.uint8, device=dev)
    kp = torch.empty((1, cap, 6), dtype=torch.float32, device=dev)
    desc = [torch.zeros((1, cap, 32), dtype=torch.uint8, device=dev) for _ in range(2)]
    count = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(2)]
    match = torch.empty((1, cap, 3), dtype=torch.int32, device=dev)
    mcount = torch.zeros(1, dtype=torch.int32, device=dev)
    h_kp = torch.empty((1, cap, 6), dtype=torch.float32).pin_memory()
    h_desc = torch.empty((1, cap, 32), dtype=torch.uint8).pin_memory()
    h_match = torch.empty((1, cap, 3), dtype=torch.int32).pin_memory()
    h_cnt = torch.empty(2, dtype=torch.int32).pin_memory()
    # a side stream: vslam_frontend_run replays a captured CUDA graph for small batches once it has seen an argument
    # set (two here: the ping-pong roles of the descriptor buffers); the legacy default stream cannot be captured,
    # so the same loop on it is the launch-by-launch comparison ("c_abi_one_frame_no_graph")
    side = torch.cuda.Stream()

    def one_frame(i, ev=None, s=side):
        with torch.cuda.stream(s):
            return _one_frame(i, ev, s)

    def _one_frame(i, ev, s):
        sptr = C.c_void_p(s.cuda_stream)
        cur, prv = i & 1, (i & 1) ^ 1
        d_frame[0].copy_(pin_in[i % 32], non_blocking=True)
        if ev:
            ev[0].record()
        _native.check(L.vslam_frontend_run(orb.handle, d_frame.data_ptr(), 1, w, desc[prv].data_ptr(), count[prv].data_ptr(),
                                           BN.RATIO, kp.data_ptr(), desc[cur].data_ptr(), count[cur].data_ptr(),
                                           match.data_ptr(), mcount.data_ptr(), sptr))
        if ev:
            ev[1].record()
        h_kp.copy_(kp, non_blocking=True)
        h_desc.copy_(desc[cur], non_blocking=True)
        h_match.copy_(match, non_blocking=True)
        h_cnt[0:1].copy_(count[cur], non_blocking=True)
        h_cnt[1:2].copy_(mcount, non_blocking=True)
        s.synchronize()

    for i in range(20):
        one_frame(i)
    wall, kern = [], []
    for i in range(n_calls):
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        t0 = time.perf_counter()
        one_frame(20 + i, ev)
        wall.append((time.perf_counter() - t0) * 1e6)
        kern.append(ev[0].elapsed_time(ev[1]) * 1e3)
    out["c_abi_one_frame"] = dict(_pcts(wall), kernel=_pcts(kern), keypoints=int(h_cnt[0]), matches=int(h_cnt[1]),
                                  call="H2D frame + vslam_frontend_run(batch 1, prev descriptors on device; CUDA graph "
                                       "replay) + D2H keypoints / descriptors / matches + sync")
    legacy = torch.cuda.default_stream()
    torch.cuda.synchronize()
    for i in range(20):
        one_frame(i, None, legacy)
    wall, kern = [], []
    for i in range(n_calls):
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        t0 = time.perf_counter()
        one_frame(20 + i, ev, legacy)
        wall.append((time.perf_counter() - t0) * 1e6)
        kern.append(ev[0].elapsed_time(ev[1]) * 1e3)
    out["c_abi_one_frame_no_graph"] = dict(_pcts(wall), kernel=_pcts(kern), keypoints=int(h_cnt[0]), matches=int(h_cnt[1]),
                                           call="the same on the legacy default stream: ~18 launches per call")
    orb.close()

    # ---- (ii) object API -------------------------------------------------------------------------------------------
    ex, mt = ORBExtractor(n_features=nf), DescriptorMatcher()
    kps_prev, d_prev = ex.detect_and_compute(frames[0])
    for i in range(5):
        _, d = ex.detect_and_compute(frames[i % 32])
        mt.match(d_prev, d)
    t_ext, t_match = [], []
    for i in range(n_calls):
        t0 = time.perf_counter()
        kps, d = ex.detect_and_compute(frames[(i + 1) % 32])
        t1 = time.perf_counter()
        m = mt.match(d_prev, d)
        t2 = time.perf_counter()
        d_prev = d
        t_ext.append((t1 - t0) * 1e6)
        t_match.append((t2 - t1) * 1e6)
    out["object_api"] = {"detect_and_compute": _pcts(t_ext), "match": _pcts(t_match),
                         "per_frame": _pcts(np.array(t_ext) + np.array(t_match)), "matches": len(m)}

    # ---- (iii) distribute_keypoints --------------------------------------------------------------------------------
    for i in range(3):
        ex.distribute_keypoints(frames[i])
    t_dist = []
    for i in range(n_calls):
        t0 = time.perf_counter()
        kps, d = ex.distribute_keypoints(frames[i % 32])
        t_dist.append((time.perf_counter() - t0) * 1e6)
    out["distribute_keypoints"] = dict(_pcts(t_dist), keypoints=len(kps), descriptors=0 if d is None else len(d))

    # ---- the reference's cv2 calls for the same work, same box -------------------------------------------------------
    try:
        import cv2
        cores = os.cpu_count() or 1
        cv2.setNumThreads(cores)
        orb_cv = cv2.ORB_create(nfeatures=nf, scaleFactor=BN.SCALE, nlevels=BN.N_LEVELS, edgeThreshold=31, firstLevel=0,
                                WTA_K=2, scoreType=cv2.ORB_HARRIS_SCORE, patchSize=31, fastThreshold=BN.FAST_THR)
        bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False)
        _, dp = orb_cv.detectAndCompute(frames[0], None)
        t_ext, t_match = [], []
        for i in range(30):
            t0 = time.perf_counter()
            _, d = orb_cv.detectAndCompute(frames[(i + 1) % 32], None)
            t1 = time.perf_counter()
            good = [p[0] for p in bf.knnMatch(dp, d, k=2) if len(p) >= 2 and p[0].distance < BN.RATIO * p[1].distance]
            t2 = time.perf_counter()
            dp = d
            if i >= 3:
                t_ext.append((t1 - t0) * 1e6)
                t_match.append((t2 - t1) * 1e6)
        out["cv2_reference"] = {"threads": cv2.getNumThreads(), "version": cv2.__version__,
                                "detect_and_compute": _pcts(t_ext), "match": _pcts(t_match),
                                "per_frame": _pcts(np.array(t_ext) + np.array(t_match))}
        # distribute_keypoints as the reference does it (extractor.py:85-144): 64 goodFeaturesToTrack calls + compute
        def ref_distribute(img):
            hh, ww = img.shape
            ch, cw = hh // 8, ww // 8
            kps = []
            for i in range(8):
                for j in range(8):
                    mask = np.zeros_like(img)
                    mask[i * ch:(i + 1) * ch, j * cw:(j + 1) * cw] = 255
                    c = cv2.goodFeaturesToTrack(img, maxCorners=nf // 64, qualityLevel=0.01, minDistance=10, mask=mask)
                    if c is not None:
                        kps += [cv2.KeyPoint(float(x), float(y), 31) for x, y in c.reshape(-1, 2)]
            return orb_cv.compute(img, kps)
        t_d = []
        for i in range(6):
            t0 = time.perf_counter()
            ref_distribute(frames[i])
            if i >= 1:
                t_d.append((time.perf_counter() - t0) * 1e6)
        out["cv2_reference"]["distribute_keypoints"] = _pcts(t_d)
    except ImportError:
        out["cv2_reference"] = {"unavailable": "cv2 is not importable on this box"}
    BN.emit(out)
